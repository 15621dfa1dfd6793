// C-ABI implementation of include/cnf2_engine.h: host-side tables (the port of what fixtrees,
// cnF2freq.cpp:3099-3187, leaves behind), device memory, launches.  No CPU compute path exists here:
// every entry point that would compute needs the CUDA device.
#include "../../include/cnf2_engine.h"
#include "cnf2_warp_kernel.cuh"
#include "cnf2_mstep_kernel.cuh"

#include <algorithm>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <map>
#include <mutex>
#include <set>
#include <string>
#include <vector>

using namespace cnf2;

namespace
{
std::string g_create_error;

struct Run  // consecutive rows holding consecutive individual ids: one 2-D copy
{
	int32_t row0, id0, n;
};

constexpr int NSLOT = 4;  // most E-steps of consecutive chromosomes in flight (the tail of one launch overlaps the next ones)

template <typename T>
struct DevBuf
{
	T* p = nullptr;
	size_t n = 0;
	cudaError_t ensure(size_t want)
	{
		if (want <= n && p) return cudaSuccess;
		if (p) cudaFree(p);
		p = nullptr;
		n = 0;
		cudaError_t e = cudaMalloc(&p, std::max<size_t>(want, 1) * sizeof(T));
		if (e == cudaSuccess) n = want;
		return e;
	}
	void release()
	{
		if (p) cudaFree(p);
		p = nullptr;
		n = 0;
	}
};
// Everything one in-flight E-step owns: its stream, scratch, statistics and the events around it.
struct Slot
{
	cudaStream_t stream = nullptr;
	cudaEvent_t ev0 = nullptr, ev1 = nullptr, done = nullptr, tail = nullptr;  // kernels start / end, all-reduce done, mark join
	DevBuf<double> alpha, qstash, ckalpha, stats, loglik;  // warp kernel: alpha ring, sweep-table ring, checkpoints
	DevBuf<double> lit_alpha, lit_fac;                       // literal kernel: alpha and its log-factor per marker
	DevBuf<int32_t> kexp, ckexp, literal;
	DevBuf<unsigned int> counter;
	int chrom = -1;
	int n_shared = 0;
	bool timed = false;  // ev0 / ev1 hold a launch whose duration has not been added to the engine's totals yet
	double* stats_shared() { return stats.p; }
	double* stats_local(int Mc) { return stats.p + (size_t)n_shared * Mc * STATS_PER_CELL; }
};
}  // namespace

struct cnf2_engine
{
	cnf2_settings settings{};
	std::string error;
	cudaStream_t stream = nullptr;  // copies of parameters, marks, microbenchmarks
	cudaStream_t comm = nullptr;
	cudaStream_t copy = nullptr;   // cnf2_get_params_packed_async: read-backs beside the kernels
	cudaStream_t comm2 = nullptr;  // control stream: the small collectives of the M-step passes (hitnnn, phase-inversion minima)    // the all-reduce hook runs here, in call order
	cudaStream_t side = nullptr;    // M-step: the relskew chains of a pass run here, beside processinfprobs
	cudaEvent_t ms_fork = nullptr, ms_join = nullptr;
	// M-step passes and phase inversions update parameters of several chromosomes from different streams: each waits for
	// the one before it through this event, and so does everything that reads parameters afterwards
	cudaEvent_t ms_chain = nullptr;
	bool chain_set = false;

	Slot slots[NSLOT];
	int next_slot = 0, last_slot = -1;
	int n_slots = 2;  // E-steps in flight (CNF2_SLOTS, 1..NSLOT)
	int n_sm = 0;

	// map
	std::vector<double> markerpos;
	std::vector<uint32_t> chromstarts;
	double genrec[3] = { 0, 0, 0 };
	int max_mc = 0;

	// pedigree
	int32_t n_ind = 0;
	std::vector<int32_t> shared_ids;
	std::vector<int32_t> rowof;   // individual id -> row, -1 = not touched by this engine
	std::vector<int32_t> rowid;   // row -> individual id
	std::vector<Run> runs;
	int32_t n_rows = 0, n_shared = 0;
	std::vector<Unit> units;

	// device
	DevBuf<uchar2> d_alleles;
	DevBuf<double2> d_sure;
	DevBuf<double> d_hw, d_rec;
	int n_ctas_warp = 0;
	int cap0 = 64, cap1 = 64;  // posterior table capacities of the warp kernel for this pedigree
	int cls = 0;               // CLS when some unit's rule-2 mask is all ones (class sums in use)
	int n_ctas_literal = 0;
	bool literal_only = false;  // debugging / A-B switch: CNF2_LITERAL=1 runs every unit through the literal kernel; always under SELFING
	bool selfing = false;       // cnf2_settings.selfing (settings.h:14)
	int n_states = NUMTYPES;    // 64, or 256 under SELFING (settings.h:27)
	size_t literal_smem = 0;    // dynamic shared memory of the literal kernel's instantiation
	DevBuf<double> d_dist;      // SELFING: distance to the next marker, laid out like the columns of d_rec's first half
	size_t pre_literal_cells = 0;  // cnf2_preprocess: (individual, marker) cells whose addvariance sums the literal kernel evaluated
	int seg_override = 0;       // CNF2_SEG: cap on the markers per segment of the warp kernel's alpha ring (experiments, tests)
	size_t ring_budget = 0;     // CNF2_RING_MB: bytes the ring of one in-flight E-step may take (0: chosen at the first E-step)
	size_t ring_budget_auto = 0;
	DevBuf<Unit> d_units;
	// M-step (row a6): inputs the E-step does not need, carried haplobase/haplocount, scratch
	DevBuf<double> d_relhaplo, d_carry_base, d_carry_count, d_ratio, d_fw;
	DevBuf<double2> d_priorsure;
	DevBuf<uchar2> d_prioral;
	DevBuf<MstepRow> d_msrows;
	DevBuf<uint8_t> d_anyinfo;
	DevBuf<EmCtl> d_ctl;
	DevBuf<double> d_minval;
	DevBuf<int32_t> d_minstart, d_invrows, d_invstart;
	std::vector<uint8_t> carry_valid;
	int32_t n_dous_total = 0;
	// phase-inversion candidates of the iteration (negshiftcands, :5261; host logic of :6270-6340)
	std::vector<int32_t> parents;                // [n_ind][2]
	std::vector<std::vector<int32_t>> kids;      // analysed children (individ::kids, :5246-5259)
	std::vector<std::vector<int32_t>> allkids;   // every individual naming this one as a parent (inverse of the pointer chains)
	std::vector<std::map<int32_t, std::pair<double, int32_t>>> candmap;  // per chromosome: id -> (val, start) of its candidate
	std::vector<uint8_t> hasmin;                 // per individual: some chromosome of this iteration holds an inversion score for it
	struct Cand { int32_t id; double val; int32_t start; bool operator<(const Cand& o) const { return id != o.id ? id < o.id : val != o.val ? val < o.val : start < o.start; } };
	std::vector<std::set<Cand>> cands;           // per chromosome
	std::vector<std::vector<double>> minval_c;   // per chromosome: minima of this iteration by individual id (empty: none yet)
	std::vector<std::vector<int32_t>> minstart_c;
	DevBuf<double> d_gmin;                       // [n_ind][2] exchange buffer between ranks
	bool have_ms_inputs = false;
	std::vector<double> h_loglik;

	// parameters stay resident for every chromosome when they fit (resident), else one chromosome at a time
	bool resident = false;
	size_t ld = 0;
	std::vector<uint8_t> have_params;
	cudaEvent_t marks[4] = { nullptr, nullptr, nullptr, nullptr };
	cudaEvent_t params_ready = nullptr;  // behind the last cnf2_set_params copies on `stream`
	int last_launches = 0;
	double estep_ms_sum = 0;  // device time of the E-step kernels (ev0 -> ev1 of every cnf2_estep), see cnf2_estep_times
	int64_t estep_ms_n = 0;
	int64_t n_kernels = 0;    // kernels launched by this engine since the last reset
	cnf2_allreduce_fn allreduce = nullptr;
	void* allreduce_ctx = nullptr;

	int fail(int code, const std::string& msg)
	{
		error = msg;
		return code;
	}
	int cuda(cudaError_t e, const char* what)
	{
		if (e == cudaSuccess) return 0;
		return fail(CNF2_ERR_CUDA, std::string(what) + ": " + cudaGetErrorString(e));
	}
	int mc(int chrom) const { return (int)(chromstarts[chrom + 1] - chromstarts[chrom]); }
	int collect_time(Slot& sl)
	{
		if (!sl.timed) return 0;
		float ms = 0;
		if (int rc = cuda(cudaEventSynchronize(sl.ev1), "cudaEventSynchronize(ev1)")) return rc;
		if (int rc = cuda(cudaEventElapsedTime(&ms, sl.ev0, sl.ev1), "cudaEventElapsedTime")) return rc;
		estep_ms_sum += ms;
		estep_ms_n++;
		sl.timed = false;
		return 0;
	}
	int wait_chain(cudaStream_t st)
	{
		if (!chain_set) return 0;
		return cuda(cudaStreamWaitEvent(st, ms_chain, 0), "cudaStreamWaitEvent(ms_chain)");
	}
	int mark_chain(cudaStream_t st)
	{
		chain_set = true;
		return cuda(cudaEventRecord(ms_chain, st), "cudaEventRecord(ms_chain)");
	}
};

#define CU(call)                                      \
	do                                                \
	{                                                 \
		int rc_ = e->cuda((call), #call);             \
		if (rc_) return rc_;                          \
	} while (0)

extern "C" void cnf2_default_settings(cnf2_settings* s)
{
	if (!s) return;
	std::memset(s, 0, sizeof(*s));
	s->numgen = 3;         // settings.h:19
	s->selfing = 0;        // settings.h:14
	s->relskewstates = 0;  // settings.h:16
	s->haplotyping = 1;    // settings.h:36
	s->doinfprobs = 1;     // settings.h:12
	s->device = 0;
	s->ctas_per_sm = 0;
}

extern "C" const char* cnf2_last_error(const cnf2_engine* e)
{
	return e ? e->error.c_str() : g_create_error.c_str();
}

extern "C" int cnf2_create(cnf2_engine** out, const cnf2_settings* settings)
{
	if (!out || !settings)
	{
		g_create_error = "cnf2_create: null argument";
		return CNF2_ERR_BAD_ARG;
	}
	*out = nullptr;
	if (settings->numgen != 3 || (settings->selfing != 0 && settings->selfing != 1) || settings->relskewstates ||
	    !settings->haplotyping || !settings->doinfprobs)
	{
		g_create_error =
		    "cnf2_create: kernels are built for NUMGEN 3, SELFING false or true, RELSKEWSTATES false, HAPLOTYPING true, "
		    "DOINFPROBS true";
		return CNF2_ERR_UNSUPPORTED;
	}
	int ndev = 0;
	cudaError_t ce = cudaGetDeviceCount(&ndev);
	if (ce != cudaSuccess || ndev <= 0 || settings->device < 0 || settings->device >= ndev)
	{
		g_create_error = std::string("cnf2_create: no usable CUDA device (") +
		                 (ce != cudaSuccess ? cudaGetErrorString(ce) : "device ordinal out of range") +
		                 "); this engine has no CPU path";
		return CNF2_ERR_NO_DEVICE;
	}
	cnf2_engine* e = new cnf2_engine;
	e->settings = *settings;
	auto bail = [&](cudaError_t err, const char* what) {
		g_create_error = std::string("cnf2_create: ") + what + ": " + cudaGetErrorString(err);
		delete e;
		return (int)CNF2_ERR_CUDA;
	};
	if ((ce = cudaSetDevice(settings->device)) != cudaSuccess) return bail(ce, "cudaSetDevice");
	cudaDeviceProp prop;
	if ((ce = cudaGetDeviceProperties(&prop, settings->device)) != cudaSuccess) return bail(ce, "cudaGetDeviceProperties");
	e->n_sm = prop.multiProcessorCount;
	if ((ce = cudaStreamCreateWithFlags(&e->stream, cudaStreamNonBlocking)) != cudaSuccess) return bail(ce, "cudaStreamCreate");
	if ((ce = cudaStreamCreateWithFlags(&e->comm, cudaStreamNonBlocking)) != cudaSuccess) return bail(ce, "cudaStreamCreate");
	if ((ce = cudaStreamCreateWithFlags(&e->comm2, cudaStreamNonBlocking)) != cudaSuccess) return bail(ce, "cudaStreamCreate");
	if ((ce = cudaStreamCreateWithFlags(&e->copy, cudaStreamNonBlocking)) != cudaSuccess) return bail(ce, "cudaStreamCreate");
	if ((ce = cudaEventCreateWithFlags(&e->params_ready, cudaEventDisableTiming)) != cudaSuccess) return bail(ce, "cudaEventCreate");
	{
		int lo_prio = 0, hi_prio = 0;
		cudaDeviceGetStreamPriorityRange(&lo_prio, &hi_prio);
		if ((ce = cudaStreamCreateWithPriority(&e->side, cudaStreamNonBlocking, hi_prio)) != cudaSuccess) return bail(ce, "cudaStreamCreate");
	}
	if ((ce = cudaEventCreateWithFlags(&e->ms_fork, cudaEventDisableTiming)) != cudaSuccess) return bail(ce, "cudaEventCreate");
	if ((ce = cudaEventCreateWithFlags(&e->ms_join, cudaEventDisableTiming)) != cudaSuccess) return bail(ce, "cudaEventCreate");
	if ((ce = cudaEventCreateWithFlags(&e->ms_chain, cudaEventDisableTiming)) != cudaSuccess) return bail(ce, "cudaEventCreate");
	for (Slot& sl : e->slots)
	{
		if ((ce = cudaStreamCreateWithFlags(&sl.stream, cudaStreamNonBlocking)) != cudaSuccess) return bail(ce, "cudaStreamCreate");
		if ((ce = cudaEventCreate(&sl.ev0)) != cudaSuccess) return bail(ce, "cudaEventCreate");
		if ((ce = cudaEventCreate(&sl.ev1)) != cudaSuccess) return bail(ce, "cudaEventCreate");
		if ((ce = cudaEventCreateWithFlags(&sl.done, cudaEventDisableTiming)) != cudaSuccess) return bail(ce, "cudaEventCreate");
		if ((ce = cudaEventCreateWithFlags(&sl.tail, cudaEventDisableTiming)) != cudaSuccess) return bail(ce, "cudaEventCreate");
	}
	// SELFING = true (row g1): 256 states; only the literal kernel is instantiated for it, every unit goes through it
	e->selfing = settings->selfing != 0;
	e->n_states = e->selfing ? NUMTYPES_SELFING : NUMTYPES;
	e->literal_smem = e->selfing ? sizeof(SmemSelfing) : sizeof(Smem);
	const void* const literal_kernel = e->selfing ? (const void*)estep_kernel<true> : (const void*)estep_kernel<false>;
	if ((ce = cudaFuncSetAttribute(literal_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)e->literal_smem)) != cudaSuccess)
		return bail(ce, "cudaFuncSetAttribute(estep_kernel)");
	int per_sm = 0;
	if ((ce = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, literal_kernel, NTHREADS, e->literal_smem)) != cudaSuccess)
		return bail(ce, "occupancy query");
	if (per_sm < 1) per_sm = 1;
	if (settings->ctas_per_sm > 0) per_sm = std::min(per_sm, settings->ctas_per_sm);
	e->n_ctas_literal = per_sm * e->n_sm;
	for (auto kern : { warp_estep_kernel<false>, warp_estep_kernel<true> })
	{
		if ((ce = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)warp_smem_bytes(64, 64, CLS))) != cudaSuccess)
			return bail(ce, "cudaFuncSetAttribute(warp_estep_kernel)");
		if ((ce = cudaFuncSetAttribute(kern, cudaFuncAttributePreferredSharedMemoryCarveout, 100)) != cudaSuccess)
			return bail(ce, "cudaFuncSetAttribute(warp_estep_kernel carveout)");
	}
	const char* lit = getenv("CNF2_LITERAL");
	e->literal_only = (lit && lit[0] == '1') || e->selfing;
	if (const char* cp = getenv("CNF2_CTAS_PER_SM")) e->settings.ctas_per_sm = atoi(cp);
	e->seg_override = settings->segment;
	if (const char* cp = getenv("CNF2_SEG")) e->seg_override = atoi(cp);
	if (const char* cp = getenv("CNF2_RING_MB")) e->ring_budget = (size_t)atoll(cp) << 20;
	if (settings->in_flight > 0) e->n_slots = std::min(NSLOT, (int)settings->in_flight);
	if (const char* ns = getenv("CNF2_SLOTS")) e->n_slots = std::max(1, std::min(NSLOT, atoi(ns)));
	*out = e;
	return CNF2_OK;
}

extern "C" void cnf2_destroy(cnf2_engine* e)
{
	if (!e) return;
	cudaSetDevice(e->settings.device);
	if (e->stream) cudaStreamSynchronize(e->stream);
#ifdef CNF2_MS_HIST
	{
		unsigned long long h[2][64];
		cudaMemcpyFromSymbol(h, ms_hist, sizeof(h));
		for (int k = 0; k < 2; k++)
		{
			fprintf(stderr, "ms_hist mode %d:", k);
			for (int i = 0; i < 64; i++) fprintf(stderr, " %llu", h[k][i]);
			fprintf(stderr, "\n");
		}
	}
#endif
	e->d_alleles.release();
	e->d_sure.release();
	e->d_hw.release();
	e->d_rec.release();
	e->d_dist.release();
	for (Slot& sl : e->slots)
	{
		if (sl.stream) cudaStreamSynchronize(sl.stream);
		sl.alpha.release(), sl.qstash.release(), sl.ckalpha.release(), sl.stats.release(), sl.loglik.release();
		sl.lit_alpha.release(), sl.lit_fac.release();
		sl.kexp.release(), sl.ckexp.release(), sl.literal.release(), sl.counter.release();
		if (sl.ev0) cudaEventDestroy(sl.ev0);
		if (sl.ev1) cudaEventDestroy(sl.ev1);
		if (sl.done) cudaEventDestroy(sl.done);
		if (sl.tail) cudaEventDestroy(sl.tail);
		if (sl.stream) cudaStreamDestroy(sl.stream);
	}
	if (e->comm) cudaStreamSynchronize(e->comm), cudaStreamDestroy(e->comm);
	if (e->comm2) cudaStreamSynchronize(e->comm2), cudaStreamDestroy(e->comm2);
	if (e->copy) cudaStreamSynchronize(e->copy), cudaStreamDestroy(e->copy);
	if (e->side) cudaStreamSynchronize(e->side), cudaStreamDestroy(e->side);
	if (e->ms_fork) cudaEventDestroy(e->ms_fork);
	if (e->ms_join) cudaEventDestroy(e->ms_join);
	if (e->ms_chain) cudaEventDestroy(e->ms_chain);
	e->d_units.release();
	e->d_relhaplo.release(), e->d_carry_base.release(), e->d_carry_count.release(), e->d_ratio.release(), e->d_fw.release();
	e->d_priorsure.release(), e->d_prioral.release(), e->d_msrows.release(), e->d_anyinfo.release(), e->d_ctl.release();
	e->d_gmin.release();
	e->d_minval.release(), e->d_minstart.release(), e->d_invrows.release(), e->d_invstart.release();
	for (cudaEvent_t m : e->marks)
		if (m) cudaEventDestroy(m);
	if (e->params_ready) cudaEventDestroy(e->params_ready);
	if (e->stream) cudaStreamDestroy(e->stream);
	delete e;
}

extern "C" int cnf2_set_map(cnf2_engine* e, const double* markerpos, int32_t n_markers, const uint32_t* chromstarts,
                            int32_t n_chrom, const double genrec[3])
{
	if (!e) return CNF2_ERR_BAD_ARG;
	if (!markerpos || !chromstarts || !genrec || n_markers <= 0 || n_chrom <= 0)
		return e->fail(CNF2_ERR_BAD_ARG, "cnf2_set_map: null or empty argument");
	if (chromstarts[0] != 0 || chromstarts[n_chrom] != (uint32_t)n_markers)
		return e->fail(CNF2_ERR_BAD_ARG, "cnf2_set_map: chromstarts must run from 0 to n_markers");
	for (int c = 0; c < n_chrom; c++)
		if (chromstarts[c + 1] <= chromstarts[c]) return e->fail(CNF2_ERR_BAD_ARG, "cnf2_set_map: empty chromosome");
	e->markerpos.assign(markerpos, markerpos + n_markers);
	e->chromstarts.assign(chromstarts, chromstarts + n_chrom + 1);
	for (int i = 0; i < 3; i++) e->genrec[i] = genrec[i];
	e->max_mc = 0;
	for (int c = 0; c < n_chrom; c++) e->max_mc = std::max(e->max_mc, e->mc(c));
	e->have_params.assign(n_chrom, 0);
	for (Slot& sl : e->slots) sl.chrom = -1;
	e->last_slot = -1;
	return CNF2_OK;
}

extern "C" int cnf2_set_shared_rows(cnf2_engine* e, const int32_t* ids, int32_t n_ids)
{
	if (!e) return CNF2_ERR_BAD_ARG;
	if (n_ids < 0 || (n_ids > 0 && !ids)) return e->fail(CNF2_ERR_BAD_ARG, "cnf2_set_shared_rows: bad list");
	e->shared_ids.assign(ids, ids + n_ids);
	return CNF2_OK;
}

extern "C" int cnf2_set_allreduce(cnf2_engine* e, cnf2_allreduce_fn fn, void* ctx)
{
	if (!e) return CNF2_ERR_BAD_ARG;
	e->allreduce = fn;
	e->allreduce_ctx = ctx;
	return CNF2_OK;
}

extern "C" int cnf2_set_pedigree(cnf2_engine* e, int32_t n_ind, const int32_t* parents, const uint8_t* empty,
                                 const uint8_t* founder_in, const int32_t* gen, const int32_t* descendants,
                                 const int32_t* dous, int32_t n_dous, const int32_t* all_dous, int32_t n_all_dous)
{
	if (!e) return CNF2_ERR_BAD_ARG;
	if (!parents || !empty || !founder_in || !gen || !descendants || n_ind <= 0 || n_dous < 0 || (n_dous && !dous))
		return e->fail(CNF2_ERR_BAD_ARG, "cnf2_set_pedigree: null or empty argument");
	if (!all_dous)
	{
		all_dous = dous;
		n_all_dous = n_dous;
	}
	auto valid = [&](int32_t id) { return id > 0 && id < n_ind; };
	for (int i = 0; i < n_ind * 2; i++)
		if (parents[i] < 0 || parents[i] >= n_ind) return e->fail(CNF2_ERR_BAD_ARG, "cnf2_set_pedigree: parent id out of range");
	for (int j = 0; j < n_dous; j++)
		if (!valid(dous[j])) return e->fail(CNF2_ERR_BAD_ARG, "cnf2_set_pedigree: analysed id out of range");
	for (int j = 0; j < n_all_dous; j++)
		if (!valid(all_dous[j])) return e->fail(CNF2_ERR_BAD_ARG, "cnf2_set_pedigree: analysed id out of range");
	for (int32_t id : e->shared_ids)
		if (!valid(id)) return e->fail(CNF2_ERR_BAD_ARG, "cnf2_set_pedigree: shared id out of range");
	cudaSetDevice(e->settings.device);

	// fixtrees' sticky founder flag in its steady state (cnF2freq.cpp:3170-3177)
	std::vector<uint8_t> founder(founder_in, founder_in + n_ind);
	auto anylev1 = [&](int ind) {
		bool any = false;
		for (int lev1 = 0; lev1 < 2; lev1++)
		{
			const int l1 = parents[ind * 2 + lev1];
			if (!l1) continue;
			bool anypars = false;
			for (int lev2 = 0; lev2 < 2; lev2++)
			{
				const int l2 = parents[l1 * 2 + lev2];
				if (l2 && !empty[l2]) anypars = true;
			}
			if (anypars || !empty[l1]) any = true;
		}
		return any;
	};
	for (int j = 0; j < n_all_dous; j++)
		if (!anylev1(all_dous[j])) founder[all_dous[j]] = 1;

	// rows: shared individuals first (same order on every rank), then the other touched ones by id
	e->n_ind = n_ind;
	e->rowof.assign(n_ind, -1);
	e->rowid.clear();
	for (int32_t id : e->shared_ids)
	{
		if (e->rowof[id] >= 0) return e->fail(CNF2_ERR_BAD_ARG, "cnf2_set_pedigree: duplicate shared id");
		e->rowof[id] = (int32_t)e->rowid.size();
		e->rowid.push_back(id);
	}
	e->n_shared = (int32_t)e->rowid.size();
	std::vector<uint8_t> touched(n_ind, 0);
	auto slot_ids = [&](int ind, int32_t ids[NSLOTS]) {
		for (int k = 0; k < NSLOTS; k++) ids[k] = 0;
		ids[0] = ind;
		for (int lev1 = 0; lev1 < 2; lev1++)
		{
			const int l1 = parents[ind * 2 + lev1];
			if (!l1) continue;
			const int base = 1 + lev1 * 3;  // :3124
			ids[base] = l1;
			for (int lev2 = 0; lev2 < 2; lev2++) ids[base + 1 + lev2] = parents[l1 * 2 + lev2];
		}
	};
	if (e->selfing)
		for (int j = 0; j < n_dous; j++)
			if (gen[dous[j]] < 2 || gen[dous[j]] > 32)  // 1 << (gen - 2) of :2095, 2321 in an int
				return e->fail(CNF2_ERR_BAD_ARG, "cnf2_set_pedigree: SELFING needs generation 2..32 for every analysed individual");
	for (int j = 0; j < n_dous; j++)
	{
		int32_t ids[NSLOTS];
		slot_ids(dous[j], ids);
		for (int k = 0; k < NSLOTS; k++)
			if (ids[k]) touched[ids[k]] = 1;
	}
	for (int id = 1; id < n_ind; id++)
		if (touched[id] && e->rowof[id] < 0)
		{
			e->rowof[id] = (int32_t)e->rowid.size();
			e->rowid.push_back(id);
		}
	e->n_rows = (int32_t)e->rowid.size();
	e->runs.clear();
	for (int r = 0; r < e->n_rows; r++)
	{
		const bool boundary = r == e->n_shared;  // a run never straddles the two statistics blocks
		if (!e->runs.empty() && !boundary && e->runs.back().id0 + e->runs.back().n == e->rowid[r]) e->runs.back().n++;
		else e->runs.push_back(Run{ r, e->rowid[r], 1 });
	}

	// individ::children as doit rebuilds it from the analysed individuals of the whole run (:5246-5259)
	std::vector<int32_t> nchildren(n_ind, 0);
	for (int j = 0; j < n_all_dous; j++)
		for (int lev1 = 0; lev1 < 2; lev1++)
			if (const int p = parents[all_dous[j] * 2 + lev1]) nchildren[p]++;

	// focus trees (fixtrees, :3099-3187)
	e->units.assign(n_dous, Unit{});
	for (int j = 0; j < n_dous; j++)
	{
		Unit& u = e->units[j];
		const int ind = dous[j];
		int32_t ids[NSLOTS];
		slot_ids(ind, ids);
		int32_t ordered[NSLOTS] = { 0 };
		std::map<int32_t, int> relmap, relmapshift;
		ordered[0] = ind;
		relmap[ind] = 1;
		relmapshift[ind] = 1;
		int flag2ignore = 1, shiftignore = 0;
		bool any1 = false;
		for (int lev1 = 0; lev1 < 2; lev1++)
		{
			const int l1 = parents[ind * 2 + lev1];
			if (!l1) continue;
			const int flag2index = 1 + lev1 * 3;  // :3124
			const int shiftval = 2 << lev1;       // :3125
			if (!empty[l1])
			{
				flag2ignore |= 1 << flag2index;
				relmap[l1] |= 1 << flag2index;
				relmapshift[l1] |= shiftval;
				ordered[flag2index] = l1;
			}
			bool anypars = false;
			for (int lev2 = 0; lev2 < 2; lev2++)
			{
				const int l2 = parents[l1 * 2 + lev2];
				if (!l2 || empty[l2]) continue;
				flag2ignore |= 1 << (flag2index + lev2 + 1);
				relmap[l2] |= 1 << (flag2index + lev2 + 1);
				relmapshift[l2] |= 0;
				ordered[flag2index + lev2 + 1] = l2;
				anypars = true;
			}
			if (anypars) shiftignore |= shiftval;
			if (anypars || !empty[l1]) any1 = true;
		}
		if (any1) shiftignore |= 1;
		u.flag2ignore = (uint8_t)(flag2ignore ^ (NUMPATHS - 1));
		u.shiftignore = (uint8_t)(shiftignore ^ (NUMSHIFTS - 1));
		u.shiftend = (uint8_t)(gen[ind] < 2 ? 2 : NUMSHIFTS);  // :5359
		u.selfgen = e->selfing ? gen[ind] - 2 : 0;
		u.out = j;
		u.descendants = (double)descendants[ind];
		// updatenegshifts (:3662-3687): who receives a phase-inversion score from this unit, and its weight
		for (int k = 0; k < NSLOTS; k++) u.nsscale[k] = 0;
		u.nsscale[0] = 1;
		for (int lev1 = 0; lev1 < 2; lev1++)
		{
			const int sp = 1 + lev1 * 3;
			if (!ids[sp]) continue;
			u.nsscale[sp] = 1;
			if (gen[ind] < 2) continue;
			for (int lev2 = 0; lev2 < 2; lev2++)
				if (ids[sp + 1 + lev2]) u.nsscale[sp + 1 + lev2] = 1.0 / nchildren[ids[sp]];
		}
		{
			const int ign = u.flag2ignore;
			u.gfixmask = (uint8_t)((ign >> 1) & (NUMTYPES - 1));
			const int freemask = ~u.gfixmask & (NUMTYPES - 1);
			int nfree = 0;
			for (int b = 0; b < 6; b++) nfree += (freemask >> b) & 1;
			u.nfree = (uint8_t)nfree;
			for (int jx = 0; jx < NUMTYPES; jx++)
			{
				int g = 0, k = 0;
				for (int b = 0; b < 6; b++)
					if ((freemask >> b) & 1) g |= ((jx >> k++) & 1) << b;
				u.expand[jx] = (uint8_t)g;
			}
			// slots visited from a non-founder root: a parent when its pointer is set, its parents when it is not a founder
			int vis = 1;
			for (int p = 0; p < 2; p++)
			{
				const int sp = 1 + 3 * p;
				if (!ids[sp]) continue;
				vis |= 1 << sp;
				if (founder[ids[sp]]) continue;
				for (int q = 0; q < 2; q++)
					if (ids[sp + 1 + q]) vis |= 1 << (sp + 1 + q);
			}
			u.vismask = (uint8_t)vis;
		}
		std::vector<int32_t> rel;  // reltree: distinct members, empty ones included, sorted (:3179-3181)
		for (int k = 0; k < NSLOTS; k++)
			if (ids[k]) rel.push_back(ids[k]);
		std::sort(rel.begin(), rel.end());
		rel.erase(std::unique(rel.begin(), rel.end()), rel.end());
		u.nrel = (uint8_t)rel.size();
		for (int k = 0; k < NSLOTS; k++)
		{
			u.row[k] = ids[k] ? e->rowof[ids[k]] : -1;
			u.founder[k] = ids[k] ? founder[ids[k]] : 0;
			u.slotrel[k] = ids[k] ? (uint8_t)(std::lower_bound(rel.begin(), rel.end(), ids[k]) - rel.begin()) : 0xff;
			u.relrow[k] = -1;
		}
		for (size_t r = 0; r < rel.size(); r++)
		{
			u.relrow[r] = e->rowof[rel[r]];
			auto it = relmap.find(rel[r]);
			u.relfilter[r] = it == relmap.end() ? 0 : (uint8_t)it->second;
			u.relshift[r] = it == relmap.end() ? 0 : (uint8_t)relmapshift[rel[r]];
			int halvings = 0;
			for (int k = 0; k < NSLOTS; k++)
				if (ordered[k] == rel[r]) halvings++;
			u.relhalvings[r] = (uint8_t)halvings;
		}
		// batched kernel: x values passing rule 2 of ignoreflag2 (:3483-3486) and the subtree configurations in use
		for (int x = 0; x < NUMPATHS; x++)
		{
			bool ok = true;
			for (int r = 0; r < u.nrel; r++)
			{
				const int filt = u.relfilter[r], filtered = x & filt;
				if (filt && filtered && filtered != filt) ok = false;
			}
			if (!ok) continue;
			u.xmask2[x >> 5] |= 1u << (x & 31);
			for (int jx = 0; jx < (1 << u.nfree); jx++)
			{
				const int g = ((x >> 1) & u.gfixmask) | u.expand[jx];
				const int flag2 = x ^ (g << 1);
				u.cfgmask[0] |= 1ull << ((g & 7) | (((flag2 >> 1) & 7) << 3));
				u.cfgmask[1] |= 1ull << (((g >> 3) & 7) | (((flag2 >> 4) & 7) << 3));
			}
		}
	}
	// warp kernel: posterior table capacities = the largest number of subtree configurations any unit uses
	e->cap0 = e->cap1 = 1;
	e->cls = 0;
	for (const Unit& u : e->units)
	{
		// units that rule 2 does not constrain take the factorised posterior: no per-configuration tables for them
		if ((u.xmask2[0] & u.xmask2[1] & u.xmask2[2] & u.xmask2[3]) == 0xffffffffu)
		{
			e->cls = CLS;
			continue;
		}
		e->cap0 = std::max(e->cap0, __builtin_popcountll(u.cfgmask[0]));
		e->cap1 = std::max(e->cap1, __builtin_popcountll(u.cfgmask[1]));
	}
	{
		int per_sm = 0;
		CU(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, e->cls ? warp_estep_kernel<true> : warp_estep_kernel<false>, 32,
		                                                 warp_smem_bytes(e->cap0, e->cap1, e->cls)));
		if (per_sm < 1) per_sm = 1;
		if (e->settings.ctas_per_sm > 0) per_sm = std::min(per_sm, e->settings.ctas_per_sm);
		e->n_ctas_warp = per_sm * e->n_sm;
		if (getenv("CNF2_DEBUG"))
			fprintf(stderr, "cnf2: warp kernel %d warps per SM, %zu B shared memory per warp (tables %d + %d entries)\n", per_sm,
			        warp_smem_bytes(e->cap0, e->cap1, e->cls), e->cap0, e->cap1);
	}
	// E-steps still in flight read d_units: they finish before the table is replaced
	for (Slot& sl : e->slots) CU(cudaStreamSynchronize(sl.stream));
	CU(e->d_units.ensure(e->units.size()));
	if (!e->units.empty())
		CU(cudaMemcpyAsync(e->d_units.p, e->units.data(), e->units.size() * sizeof(Unit), cudaMemcpyHostToDevice, e->stream));
	for (Slot& sl : e->slots)
	{
		CU(sl.loglik.ensure(e->units.size() * 2));
		CU(sl.counter.ensure(4));
		CU(sl.literal.ensure(e->units.size()));
	}
	e->h_loglik.assign(e->units.size() * 2, 0.0);
	e->n_dous_total = n_all_dous;
	e->parents.assign(parents, parents + (size_t)n_ind * 2);
	e->kids.assign(n_ind, {});
	for (int j = 0; j < n_all_dous; j++)
		for (int lev1 = 0; lev1 < 2; lev1++)
			if (const int p = parents[all_dous[j] * 2 + lev1]) e->kids[p].push_back(all_dous[j]);
	e->allkids.assign(n_ind, {});
	for (int i = 1; i < n_ind; i++)
		for (int lev1 = 0; lev1 < 2; lev1++)
			if (const int p = parents[i * 2 + lev1])
				if (e->allkids[p].empty() || e->allkids[p].back() != i) e->allkids[p].push_back(i);
	CU(cudaStreamSynchronize(e->stream));
	std::fill(e->have_params.begin(), e->have_params.end(), 0);
	for (Slot& sl : e->slots) sl.chrom = -1;
	e->last_slot = -1;
	// 26 bytes of parameters per (row, marker): keep all chromosomes on the device when that is a small part of HBM
	size_t free_b = 0, total_b = 0;
	CU(cudaMemGetInfo(&free_b, &total_b));
	const size_t M = e->markerpos.size();
	e->resident = !e->markerpos.empty() && (double)e->n_rows * M * 26.0 < 0.4 * (double)free_b;
	e->ld = 0;  // parameters are stored per chromosome: [n_rows][Mc] blocks, chromosome c at n_rows * chromstarts[c] when resident
	return CNF2_OK;
}

static int set_params_impl(cnf2_engine* e, int32_t chrom, const uint8_t* alleles, const double* sure, const double* haploweight,
                           bool packed)
{
	if (!e) return CNF2_ERR_BAD_ARG;
	if (e->chromstarts.empty() || e->rowof.empty()) return e->fail(CNF2_ERR_STATE, "cnf2_set_params: set_map and set_pedigree first");
	if (chrom < 0 || chrom + 1 >= (int)e->chromstarts.size() || !alleles || !sure || !haploweight)
		return e->fail(CNF2_ERR_BAD_ARG, "cnf2_set_params: bad argument");
	cudaSetDevice(e->settings.device);
	if (!e->resident)  // the buffers hold one chromosome: kernels still reading the previous one finish first
		for (Slot& sl : e->slots) CU(cudaStreamSynchronize(sl.stream));
	const size_t M = e->markerpos.size();
	const size_t c0 = e->chromstarts[chrom];
	const size_t Mc = e->mc(chrom), ld = Mc;
	const size_t col0 = e->resident ? c0 : 0;
	const size_t total = (size_t)e->n_rows * (e->resident ? M : (size_t)e->max_mc), blk = (size_t)e->n_rows * col0;
	CU(e->d_alleles.ensure(total));
	CU(e->d_sure.ensure(total));
	CU(e->d_hw.ensure(total));
	if (int rc = e->wait_chain(e->stream)) return rc;
	if (packed)
	{
		// row-ordered contiguous blocks [n_rows][Mc]: one unstrided copy per array
		CU(cudaMemcpyAsync(e->d_alleles.p + blk, alleles, (size_t)e->n_rows * Mc * 2, cudaMemcpyHostToDevice, e->stream));
		CU(cudaMemcpyAsync(e->d_sure.p + blk, sure, (size_t)e->n_rows * Mc * 16, cudaMemcpyHostToDevice, e->stream));
		CU(cudaMemcpyAsync(e->d_hw.p + blk, haploweight, (size_t)e->n_rows * Mc * 8, cudaMemcpyHostToDevice, e->stream));
	}
	else
		for (const Run& r : e->runs)
		{
			const size_t src = (size_t)r.id0 * M + c0, dst = blk + (size_t)r.row0 * ld;
			CU(cudaMemcpy2DAsync(e->d_alleles.p + dst, ld * 2, alleles + src * 2, M * 2, Mc * 2, r.n, cudaMemcpyHostToDevice, e->stream));
			CU(cudaMemcpy2DAsync(e->d_sure.p + dst, ld * 16, sure + src * 2, M * 16, Mc * 16, r.n, cudaMemcpyHostToDevice, e->stream));
			CU(cudaMemcpy2DAsync(e->d_hw.p + dst, ld * 8, haploweight + src, M * 8, Mc * 8, r.n, cudaMemcpyHostToDevice, e->stream));
		}
	// recombination fractions towards the next marker (:2286; getactrec :777-780 ignores sex and marker)
	std::vector<double> rec(2 * Mc, 0.0);
	for (size_t m = 0; m + 1 < Mc; m++)
	{
		const double dist = e->markerpos[c0 + m + 1] - e->markerpos[c0 + m];
		if (!(dist > 0)) continue;  // same position: no transition (:2273)
		rec[m] = 0.5 * (1.0 - std::exp(e->genrec[0] * dist));
		rec[Mc + m] = 0.5 * (1.0 - std::exp(e->genrec[1] * dist));
	}
	CU(e->d_rec.ensure(2 * (e->resident ? M : (size_t)e->max_mc)));
	CU(cudaMemcpyAsync(e->d_rec.p + 2 * col0, rec.data(), rec.size() * sizeof(double), cudaMemcpyHostToDevice, e->stream));
	if (e->selfing)
	{
		// the selfing generations' rate depends on the unit (:2286: selfgen * genrec[2] * dist): the kernel gets the distances
		std::vector<double> dist(Mc, 0.0);
		for (size_t m = 0; m + 1 < Mc; m++) dist[m] = e->markerpos[c0 + m + 1] - e->markerpos[c0 + m];
		CU(e->d_dist.ensure(e->resident ? M : (size_t)e->max_mc));
		CU(cudaMemcpyAsync(e->d_dist.p + col0, dist.data(), dist.size() * sizeof(double), cudaMemcpyHostToDevice, e->stream));
	}
	CU(cudaEventRecord(e->params_ready, e->stream));
	// async_params: the copies from page-locked caller buffers are left in flight (the E-step waits for them on the
	// device); otherwise caller buffers may be reused after return.  rec is pageable: staged before the call returns.
	if (!e->settings.async_params) CU(cudaStreamSynchronize(e->stream));
	if (!e->resident) std::fill(e->have_params.begin(), e->have_params.end(), 0);
	e->have_params[chrom] = 1;
	return CNF2_OK;
}

extern "C" int cnf2_set_params(cnf2_engine* e, int32_t chrom, const uint8_t* alleles, const double* sure,
                               const double* haploweight)
{
	return set_params_impl(e, chrom, alleles, sure, haploweight, false);
}

extern "C" int cnf2_set_params_packed(cnf2_engine* e, int32_t chrom, const uint8_t* alleles, const double* sure,
                                      const double* haploweight)
{
	return set_params_impl(e, chrom, alleles, sure, haploweight, true);
}

static Slot* slot_of(cnf2_engine* e, int32_t chrom)
{
	for (Slot& o : e->slots)
		if (chrom >= 0 && o.chrom == chrom) return &o;
	return nullptr;
}

extern "C" int cnf2_estep(cnf2_engine* e, int32_t chrom)
{
	if (!e) return CNF2_ERR_BAD_ARG;
	if (chrom < 0 || chrom >= (int)e->have_params.size() || !e->have_params[chrom])
		return e->fail(CNF2_ERR_STATE, "cnf2_estep: cnf2_set_params for this chromosome first");
	cudaSetDevice(e->settings.device);
	const int Mc = e->mc(chrom);
	const int n_units = (int)e->units.size();
	// the next slot; what it held (the E-step before the previous one) is overwritten
	Slot& sl = e->slots[e->next_slot];
	e->last_slot = e->next_slot;
	e->next_slot = (e->next_slot + 1) % e->n_slots;
	for (Slot& o : e->slots)
		if (o.chrom == chrom) o.chrom = -1;
	// Scratch of the sweeps.  The warp kernel keeps, per resident warp, a ring of `seg` markers (alpha before each marker's
	// emission for the 8 shift modes, its exponents, the marker's sweep table: 4,672 B per marker) and one alpha checkpoint per
	// segment; a chromosome longer than `seg` is walked segment by segment, its alphas recomputed from the checkpoints
	// (see warp_estep_kernel).  seg = the whole chromosome while the ring stays within the budget, so the footprint does not
	// grow with the chromosome beyond that.  The literal kernel (a few CTAs, rare units) keeps every marker.
	const int n_warps = std::max(1, std::min(e->n_ctas_warp, n_units));
	const int n_lit = std::max(1, std::min(e->n_ctas_literal, n_units));
	const size_t ring_per_marker = (size_t)NUMSHIFTS * NUMTYPES * 8 + NUMSHIFTS * 4 + QS * 8;
	if (!e->ring_budget_auto)
	{
		// the rings of all in-flight E-steps together may take a third of what is free when the first E-step comes (parameters
		// and M-step inputs are resident by then): with one slot a whole chromosome of 5,000 markers fits the ring, with three
		// slots on an otherwise empty B200 each gets ~19 GB; on a nearly full device the segments get shorter instead
		size_t free_b = 0, total_b = 0;
		CU(cudaMemGetInfo(&free_b, &total_b));
		e->ring_budget_auto = std::max<size_t>((size_t)1 << 30, free_b / 3 / (size_t)e->n_slots);
	}
	auto seg_for = [&](int mc) {
		int sg = ((mc + WB - 1) / WB) * WB;
		const size_t budget = e->ring_budget ? e->ring_budget : e->ring_budget_auto;
		const size_t fit = budget / ((size_t)n_warps * ring_per_marker);
		if ((size_t)sg > fit) sg = (int)std::max<size_t>(WB, fit / WB * WB);
		if (e->seg_override > 0) sg = std::min(sg, std::max(WB, e->seg_override / WB * WB));
		return sg;
	};
	const int seg = seg_for(Mc), nseg = (Mc + seg - 1) / seg;
	if (!e->literal_only)
	{
		// sized for the longest chromosome, so that no later E-step of the run allocates (seg and nseg grow with the length)
		const int seg_max = seg_for(e->max_mc), nseg_max = (e->max_mc + seg_max - 1) / seg_max;
		CU(sl.alpha.ensure((size_t)n_warps * seg_max * NUMSHIFTS * NUMTYPES));
		CU(sl.kexp.ensure((size_t)n_warps * seg_max * NUMSHIFTS));
		CU(sl.qstash.ensure((size_t)n_warps * seg_max * QS));
		CU(sl.ckalpha.ensure((size_t)n_warps * nseg_max * NUMSHIFTS * NUMTYPES));
		CU(sl.ckexp.ensure((size_t)n_warps * nseg_max * NUMSHIFTS));
	}
	CU(sl.lit_alpha.ensure((size_t)n_lit * e->max_mc * NUMSHIFTS * e->n_states));
	CU(sl.lit_fac.ensure((size_t)n_lit * e->max_mc * NUMSHIFTS));
	CU(sl.stats.ensure((size_t)e->n_rows * e->max_mc * STATS_PER_CELL));
	sl.n_shared = e->n_shared;
	sl.chrom = chrom;
	if (int rc = e->collect_time(sl)) return rc;  // the launch this slot held before (n_slots E-steps ago): long finished
	CU(cudaStreamWaitEvent(sl.stream, e->params_ready, 0));
	if (int rc = e->wait_chain(sl.stream)) return rc;  // the parameters an earlier M-step pass or phase inversion is still writing
	CU(cudaEventRecord(sl.ev0, sl.stream));
	CU(cudaMemsetAsync(sl.stats.p, 0, (size_t)e->n_rows * Mc * STATS_PER_CELL * sizeof(double), sl.stream));
	CU(cudaMemsetAsync(sl.counter.p, 0, 4 * sizeof(unsigned int), sl.stream));
	e->last_launches = 0;
	if (n_units > 0)
	{
		WarpArgs wa{};
		EstepArgs& a = wa.e;
		const size_t col0 = e->resident ? e->chromstarts[chrom] : 0;
		a.alleles = e->d_alleles.p + (size_t)e->n_rows * col0;
		a.sure = e->d_sure.p + (size_t)e->n_rows * col0;
		a.hw = e->d_hw.p + (size_t)e->n_rows * col0;
		a.rec0 = e->d_rec.p + 2 * col0;
		a.rec1 = e->d_rec.p + 2 * col0 + Mc;
		a.units = e->d_units.p;
		a.n_units = n_units;
		a.Mc = Mc;
		a.ld = Mc;
		a.counter = sl.counter.p;
		a.alpha = sl.alpha.p;
		a.alphafac = nullptr;
		a.stats_shared = sl.stats_shared();
		a.stats_local = sl.stats_local(Mc);
		a.n_shared = e->n_shared;
		a.n_local = e->n_rows - e->n_shared;
		a.loglik = sl.loglik.p;
		a.unit_list = nullptr;
		a.list_count = nullptr;
		if (!e->literal_only)
		{
			// one warp per unit; units it cannot take (founder roots, out-of-range parameters) are listed for the literal kernel
			wa.kexp = sl.kexp.p;
			wa.qstash = sl.qstash.p;
			wa.ckalpha = sl.ckalpha.p;
			wa.ckexp = sl.ckexp.p;
			wa.seg = seg;
			wa.nseg = nseg;
			wa.turns = 1;
			wa.literal_units = sl.literal.p;
			wa.literal_count = sl.counter.p + 1;
			wa.cap0 = e->cap0;
			wa.cap1 = e->cap1;
			wa.cls = e->cls;
			const int grid = std::max(1, std::min(e->n_ctas_warp, n_units));
			const size_t smem = warp_smem_bytes(e->cap0, e->cap1, e->cls);
			if (e->cls) warp_estep_kernel<true><<<grid, 32, smem, sl.stream>>>(wa);
			else warp_estep_kernel<false><<<grid, 32, smem, sl.stream>>>(wa);
			e->n_kernels++;
			CU(cudaGetLastError());
			e->last_launches++;
			a.unit_list = sl.literal.p;
			a.list_count = sl.counter.p + 1;
			a.counter = sl.counter.p + 2;
		}
		a.alpha = sl.lit_alpha.p;
		a.alphafac = sl.lit_fac.p;
		const int lit_grid = std::max(1, std::min(e->n_ctas_literal, n_units));
		if (e->selfing)
		{
			a.dist = e->d_dist.p + col0;
			a.genrec2 = e->genrec[2];
			estep_kernel<true><<<lit_grid, NTHREADS, e->literal_smem, sl.stream>>>(a);
		}
		else estep_kernel<false><<<lit_grid, NTHREADS, e->literal_smem, sl.stream>>>(a);
		e->n_kernels++;
		CU(cudaGetLastError());
		e->last_launches++;
	}
	CU(cudaEventRecord(sl.ev1, sl.stream));
	sl.timed = true;
	if (e->allreduce && e->n_shared > 0)
	{
		// the hook runs on one stream for all slots, so every rank issues its collectives in the order of its cnf2_estep calls
		CU(cudaStreamWaitEvent(e->comm, sl.ev1, 0));
		const int rc = e->allreduce(e->allreduce_ctx, sl.stats_shared(), (size_t)e->n_shared * Mc * STATS_PER_CELL, (void*)e->comm);
		if (rc) return e->fail(CNF2_ERR_CUDA, "cnf2_estep: all-reduce hook failed");
		CU(cudaEventRecord(sl.done, e->comm));
		CU(cudaStreamWaitEvent(sl.stream, sl.done, 0));
	}
	return CNF2_OK;
}

extern "C" int cnf2_sync(cnf2_engine* e)
{
	if (!e) return CNF2_ERR_BAD_ARG;
	cudaSetDevice(e->settings.device);
	for (Slot& sl : e->slots) CU(cudaStreamSynchronize(sl.stream));
	CU(cudaStreamSynchronize(e->comm));
	CU(cudaStreamSynchronize(e->comm2));
	CU(cudaStreamSynchronize(e->side));
	CU(cudaStreamSynchronize(e->stream));
	CU(cudaStreamSynchronize(e->copy));
	CU(cudaGetLastError());
	return CNF2_OK;
}

extern "C" int cnf2_get_stats(cnf2_engine* e, int32_t chrom, double* haplobase, double* haplocount, double* infprobs,
                              double* homozyg, double* loglik)
{
	if (!e) return CNF2_ERR_BAD_ARG;
	Slot* slp = nullptr;
	for (Slot& o : e->slots)
		if (chrom >= 0 && o.chrom == chrom) slp = &o;
	if (!slp) return e->fail(CNF2_ERR_STATE, "cnf2_get_stats: cnf2_estep for this chromosome first (only the last two E-steps are kept)");
	Slot& sl = *slp;
	cudaSetDevice(e->settings.device);
	const size_t M = e->markerpos.size(), c0 = e->chromstarts[chrom], Mc = e->mc(chrom);
	for (const Run& r : e->runs)
	{
		const bool sh = r.row0 < e->n_shared;
		const StatsCell src = sh ? stats_cell(sl.stats_shared(), e->n_shared, Mc, r.row0, 0)
		                         : stats_cell(sl.stats_local((int)Mc), e->n_rows - e->n_shared, Mc, r.row0 - e->n_shared, 0);
		const size_t dst = (size_t)r.id0 * M + c0;
		if (haplobase)
			CU(cudaMemcpy2DAsync(haplobase + dst, M * 8, src.haplobase, Mc * 8, Mc * 8, r.n, cudaMemcpyDeviceToHost, sl.stream));
		if (haplocount)
			CU(cudaMemcpy2DAsync(haplocount + dst, M * 8, src.haplocount, Mc * 8, Mc * 8, r.n, cudaMemcpyDeviceToHost, sl.stream));
		if (infprobs)
			CU(cudaMemcpy2DAsync(infprobs + dst * 4, M * 32, src.infprobs, Mc * 32, Mc * 32, r.n, cudaMemcpyDeviceToHost, sl.stream));
		if (homozyg)
			CU(cudaMemcpy2DAsync(homozyg + dst * 2, M * 16, src.homozyg, Mc * 16, Mc * 16, r.n, cudaMemcpyDeviceToHost, sl.stream));
	}
	if (loglik && !e->units.empty())
		CU(cudaMemcpyAsync(e->h_loglik.data(), sl.loglik.p, e->units.size() * 2 * sizeof(double), cudaMemcpyDeviceToHost, sl.stream));
	CU(cudaStreamSynchronize(sl.stream));
	CU(cudaGetLastError());
	if (loglik)
	{
		const size_t C = e->chromstarts.size() - 1;
		for (size_t j = 0; j < e->units.size(); j++)
		{
			loglik[(j * C + chrom) * 2 + 0] = e->h_loglik[j * 2 + 0];
			loglik[(j * C + chrom) * 2 + 1] = e->h_loglik[j * 2 + 1];
		}
	}
	return CNF2_OK;
}

// device arrays of the M-step: every chromosome's block, the control state
static int ms_alloc(cnf2_engine* e)
{
	if (e->chromstarts.empty() || e->rowof.empty()) return e->fail(CNF2_ERR_STATE, "cnf2_set_mstep_inputs: set_map and set_pedigree first");
	if (!e->resident) return e->fail(CNF2_ERR_STATE, "cnf2_set_mstep_inputs: the M-step needs the parameters of all chromosomes resident");
	cudaSetDevice(e->settings.device);
	const size_t M = e->markerpos.size(), total = (size_t)e->n_rows * M;
	CU(e->d_relhaplo.ensure(total));
	CU(e->d_prioral.ensure(total));
	CU(e->d_priorsure.ensure(total));
	CU(e->d_carry_base.ensure(total));
	CU(e->d_carry_count.ensure(total));
	CU(e->d_ratio.ensure((size_t)e->n_rows * e->markerpos.size()));  // per chromosome block: a pass computes them all before it uses them
	CU(e->d_fw.ensure(total * 2));  // per chromosome: the chains of every chromosome of a pass run in one launch
	CU(e->d_anyinfo.ensure((size_t)e->n_rows * (e->chromstarts.size() - 1)));
	if (!e->d_ctl.p)
	{
		CU(e->d_ctl.ensure(1));
		EmCtl init{};
		init.scalefactor = 0.013;  // :3573
		init.entropyfactor = 1;    // :3574
		init.n_dous_total = e->n_dous_total;
		CU(cudaMemcpyAsync(e->d_ctl.p, &init, sizeof(init), cudaMemcpyHostToDevice, e->stream));
	}
	CU(e->d_minval.ensure(e->n_rows));
	CU(e->d_minstart.ensure(e->n_rows));
	CU(e->d_msrows.ensure(e->n_rows));
	return CNF2_OK;
}

extern "C" int cnf2_set_mstep_rows(cnf2_engine* e, const uint8_t* hasprior, const int32_t* children, const uint8_t* empty,
                                   const int32_t* descendants)
{
	if (!e || !hasprior || !children || !empty || !descendants) return CNF2_ERR_BAD_ARG;
	if (int rc = ms_alloc(e)) return rc;
	std::vector<MstepRow> rows(e->n_rows);
	for (int32_t r = 0; r < e->n_rows; r++)
	{
		const int32_t id = e->rowid[r];
		rows[r] = MstepRow{ children[id], descendants[id], empty[id], hasprior[id], { 0, 0 } };
	}
	CU(cudaMemcpyAsync(e->d_msrows.p, rows.data(), rows.size() * sizeof(MstepRow), cudaMemcpyHostToDevice, e->stream));
	CU(cudaStreamSynchronize(e->stream));
	e->carry_valid.assign(e->chromstarts.size() - 1, 0);
	e->have_ms_inputs = true;
	return CNF2_OK;
}

extern "C" int cnf2_set_mstep_chrom_packed(cnf2_engine* e, int32_t chrom, const double* relhaplo, const uint8_t* prioralleles,
                                           const double* priorsure)
{
	if (!e || !relhaplo || !prioralleles || !priorsure) return CNF2_ERR_BAD_ARG;
	if (int rc = ms_alloc(e)) return rc;
	if (chrom < 0 || chrom + 1 >= (int)e->chromstarts.size()) return e->fail(CNF2_ERR_BAD_ARG, "cnf2_set_mstep_chrom_packed: bad chromosome");
	const size_t Mc = e->mc(chrom), n = (size_t)e->n_rows * Mc, blk = (size_t)e->n_rows * e->chromstarts[chrom];
	if (int rc = e->wait_chain(e->stream)) return rc;
	CU(cudaMemcpyAsync(e->d_relhaplo.p + blk, relhaplo, n * 8, cudaMemcpyHostToDevice, e->stream));
	CU(cudaMemcpyAsync(e->d_prioral.p + blk, prioralleles, n * 2, cudaMemcpyHostToDevice, e->stream));
	CU(cudaMemcpyAsync(e->d_priorsure.p + blk, priorsure, n * 16, cudaMemcpyHostToDevice, e->stream));
	CU(cudaEventRecord(e->params_ready, e->stream));
	if (!e->settings.async_params) CU(cudaStreamSynchronize(e->stream));
	return CNF2_OK;
}

extern "C" int cnf2_set_mstep_inputs(cnf2_engine* e, const double* relhaplo, const uint8_t* prioralleles, const double* priorsure,
                                     const uint8_t* hasprior, const int32_t* children, const uint8_t* empty,
                                     const int32_t* descendants)
{
	if (!e || !relhaplo || !prioralleles || !priorsure || !hasprior || !children || !empty || !descendants) return CNF2_ERR_BAD_ARG;
	if (int rc = ms_alloc(e)) return rc;
	const size_t M = e->markerpos.size();
	for (size_t c = 0; c + 1 < e->chromstarts.size(); c++)
	{
		const size_t c0 = e->chromstarts[c], Mc = e->mc((int)c), blk = (size_t)e->n_rows * c0;
		for (const Run& r : e->runs)
		{
			const size_t src = (size_t)r.id0 * M + c0, dst = blk + (size_t)r.row0 * Mc;
			CU(cudaMemcpy2DAsync(e->d_relhaplo.p + dst, Mc * 8, relhaplo + src, M * 8, Mc * 8, r.n, cudaMemcpyHostToDevice, e->stream));
			CU(cudaMemcpy2DAsync(e->d_prioral.p + dst, Mc * 2, prioralleles + src * 2, M * 2, Mc * 2, r.n, cudaMemcpyHostToDevice, e->stream));
			CU(cudaMemcpy2DAsync(e->d_priorsure.p + dst, Mc * 16, priorsure + src * 2, M * 16, Mc * 16, r.n, cudaMemcpyHostToDevice, e->stream));
		}
	}
	CU(cudaStreamSynchronize(e->stream));
	return cnf2_set_mstep_rows(e, hasprior, children, empty, descendants);
}

extern "C" int cnf2_begin_iteration(cnf2_engine* e)
{
	if (!e) return CNF2_ERR_BAD_ARG;
	std::fill(e->carry_valid.begin(), e->carry_valid.end(), 0);  // doit zeroes haplobase / haplocount, :5222-5234
	const size_t C = e->chromstarts.empty() ? 0 : e->chromstarts.size() - 1;
	e->cands.assign(C, {});       // negshiftcands is local to doit (:5261)
	e->candmap.assign(C, {});
	e->hasmin.assign(e->n_ind, 0);
	e->minval_c.assign(C, {});    // negshift is zeroed after every iteration's dump (:8187)
	e->minstart_c.assign(C, {});
	return CNF2_OK;
}

// One M-step pass behind chromosome `chrom`'s E-step, enqueued on that E-step's stream.  set: scalefactor / entropyfactor
// handed in (cnf2_mstep); control: run the step-size control of :6373-6390 on the device (cnf2_mstep_pass).
static int enqueue_pass(cnf2_engine* e, int32_t chrom, int any, int set, double sf, double ent, int control, Slot** out_slot)
{
	if (!e->have_ms_inputs) return e->fail(CNF2_ERR_STATE, "cnf2_mstep: cnf2_set_mstep_inputs first");
	Slot* slp = nullptr;
	for (Slot& o : e->slots)
		if (chrom >= 0 && o.chrom == chrom) slp = &o;
	if (!slp) return e->fail(CNF2_ERR_STATE, "cnf2_mstep: cnf2_estep for this chromosome first (its statistics must still be in flight)");
	Slot& sl = *slp;
	*out_slot = slp;
	cudaSetDevice(e->settings.device);
	const int C = (int)e->chromstarts.size() - 1;
	if (int rc = e->wait_chain(sl.stream)) return rc;
	ms_begin_kernel<<<1, 1, 0, sl.stream>>>(e->d_ctl.p, any, set, sf, ent);
	e->n_kernels++;
	if (e->n_rows > 0)
	{
		MstepArgs g{};
		g.alleles = e->d_alleles.p;
		g.sure = e->d_sure.p;
		g.hw = e->d_hw.p;
		g.relhaplo = e->d_relhaplo.p;
		g.prioralleles = e->d_prioral.p;
		g.priorsure = e->d_priorsure.p;
		g.rows = e->d_msrows.p;
		g.stats_shared = sl.stats_shared();
		g.stats_local = sl.stats_local(e->mc(chrom));
		g.carry_base = e->d_carry_base.p;
		g.carry_count = e->d_carry_count.p;
		g.ratio = e->d_ratio.p;
		g.fw = e->d_fw.p;
		g.anyinfo = e->d_anyinfo.p;
		g.ctl = e->d_ctl.p;
		g.n_rows = e->n_rows;
		g.n_shared = e->n_shared;
		// processinfprobs only has statistics for this chromosome; updatehaploweights also walks the chromosomes whose
		// pass came earlier in the iteration, with the haplobase / haplocount that pass left behind (:6239-6367).
		// The relskew chains (one thread per (row, chromosome), latency-bound) read haploweight and haplocount only, which
		// processinfprobs does not touch: they run on a side stream beside it.
		std::vector<int> in_pass;
		for (int c = 0; c < C; c++)
			if (c == chrom || e->carry_valid[c]) in_pass.push_back(c);
		std::vector<MstepPass> launches;
		for (size_t i = 0; i < in_pass.size(); i += MS_MAXLIST)
		{
			MstepPass p{};
			p.g = g;
			p.fresh_index = -1;
			p.n = (int)std::min<size_t>(MS_MAXLIST, in_pass.size() - i);
			for (int k = 0; k < p.n; k++)
			{
				const int c = in_pass[i + k];
				p.c0[k] = (int32_t)e->chromstarts[c];
				p.Mc[k] = e->mc(c);
				p.cidx[k] = c;
				if (c == chrom) p.fresh_index = k;
			}
			launches.push_back(p);
		}
		const unsigned gr = (unsigned)((e->n_rows + 127) / 128);
		CU(cudaEventRecord(e->ms_fork, sl.stream));
		CU(cudaStreamWaitEvent(e->side, e->ms_fork, 0));
		for (const MstepPass& p : launches) e->n_kernels++, mstep_relskew_kernel<<<dim3(gr, p.n), 128, 0, e->side>>>(p);
		CU(cudaEventRecord(e->ms_join, e->side));
		{
			MstepArgs a = g;
			const size_t off = (size_t)e->n_rows * e->chromstarts[chrom];
			a.alleles += off, a.sure += off, a.hw += off, a.relhaplo += off, a.prioralleles += off, a.priorsure += off;
			a.Mc = e->mc(chrom);
			a.fresh = 1;
			mstep_infprobs_kernel<<<(unsigned)(((size_t)e->n_rows * a.Mc + 127) / 128), 128, 0, sl.stream>>>(a);
			e->n_kernels++;
		}
		CU(cudaStreamWaitEvent(sl.stream, e->ms_join, 0));
		for (const MstepPass& p : launches)
		{
			int maxmc = 0;
			for (int k = 0; k < p.n; k++) maxmc = std::max(maxmc, p.Mc[k]);
			mstep_haploweight_kernel<<<dim3((unsigned)(((size_t)e->n_rows * maxmc + 127) / 128), p.n), 128, 0, sl.stream>>>(p);
			e->n_kernels++;
		}
		CU(cudaGetLastError());
	}
	e->carry_valid[chrom] = 1;
	int reduced = 0;
	if (control && e->allreduce && e->n_shared > 0)
	{
		// global hitnnn = sum over ranks of the local counts + the shared rows' count (identical on every rank)
		ms_pack_kernel<<<1, 1, 0, sl.stream>>>(e->d_ctl.p);
		e->n_kernels++;
		CU(cudaEventRecord(sl.tail, sl.stream));
		// on the control stream: behind the statistics all-reduces of the E-steps issued ahead (e->comm) this one-double sum
		// would wait for those E-steps to finish, and the next pass with it
		CU(cudaStreamWaitEvent(e->comm2, sl.tail, 0));
		const int rc = e->allreduce(e->allreduce_ctx, &e->d_ctl.p->hit_local_d, 1, (void*)e->comm2);
		if (rc) return e->fail(CNF2_ERR_CUDA, "cnf2_mstep_pass: all-reduce hook failed");
		CU(cudaEventRecord(sl.done, e->comm2));
		CU(cudaStreamWaitEvent(sl.stream, sl.done, 0));
		reduced = 1;
	}
	ms_end_kernel<<<1, 1, 0, sl.stream>>>(e->d_ctl.p, reduced, control);
	e->n_kernels++;
	CU(cudaGetLastError());
	return e->mark_chain(sl.stream);
}

extern "C" int cnf2_mstep(cnf2_engine* e, int32_t chrom, double scalefactor, double entropyfactor, int64_t* hitnnn,
                          int64_t* hitnnn_shared)
{
	if (!e) return CNF2_ERR_BAD_ARG;
	Slot* sl = nullptr;
	const int rc = enqueue_pass(e, chrom, 0, 1, scalefactor, entropyfactor, 0, &sl);
	if (rc) return rc;
	EmCtl c{};
	CU(cudaMemcpyAsync(&c, e->d_ctl.p, sizeof(c), cudaMemcpyDeviceToHost, sl->stream));
	CU(cudaStreamSynchronize(sl->stream));
	CU(cudaGetLastError());
	if (hitnnn) *hitnnn = (int64_t)(c.hit[0] + c.hit[1]);
	if (hitnnn_shared) *hitnnn_shared = (int64_t)c.hit[1];
	return CNF2_OK;
}

extern "C" int cnf2_mstep_pass(cnf2_engine* e, int32_t chrom, int32_t any_inversion)
{
	if (!e) return CNF2_ERR_BAD_ARG;
	Slot* sl = nullptr;
	return enqueue_pass(e, chrom, any_inversion ? 1 : 0, 0, 0.0, 0.0, 1, &sl);
}

extern "C" int cnf2_set_em_state(cnf2_engine* e, double scalefactor, double entropyfactor, int32_t oldhitnnn, int32_t oldhitnnn2)
{
	if (!e) return CNF2_ERR_BAD_ARG;
	if (!e->have_ms_inputs) return e->fail(CNF2_ERR_STATE, "cnf2_set_em_state: cnf2_set_mstep_inputs first");
	cudaSetDevice(e->settings.device);
	CU(cnf2_sync(e) == CNF2_OK ? cudaSuccess : cudaErrorUnknown);
	EmCtl c{};
	c.scalefactor = scalefactor;
	c.entropyfactor = entropyfactor;
	c.oldhitnnn = oldhitnnn;
	c.oldhitnnn2 = oldhitnnn2;
	c.n_dous_total = e->n_dous_total;
	CU(cudaMemcpyAsync(e->d_ctl.p, &c, sizeof(c), cudaMemcpyHostToDevice, e->stream));
	CU(cudaStreamSynchronize(e->stream));
	return CNF2_OK;
}

extern "C" int cnf2_get_em_state(cnf2_engine* e, double* scalefactor, double* entropyfactor, int32_t* oldhitnnn, int32_t* oldhitnnn2,
                                 int64_t* hitnnn_last)
{
	if (!e) return CNF2_ERR_BAD_ARG;
	if (!e->have_ms_inputs) return e->fail(CNF2_ERR_STATE, "cnf2_get_em_state: cnf2_set_mstep_inputs first");
	cudaSetDevice(e->settings.device);
	{
		const int rc = cnf2_sync(e);
		if (rc) return rc;
	}
	EmCtl c{};
	CU(cudaMemcpy(&c, e->d_ctl.p, sizeof(c), cudaMemcpyDeviceToHost));
	if (scalefactor) *scalefactor = c.scalefactor;
	if (entropyfactor) *entropyfactor = c.entropyfactor;
	if (oldhitnnn) *oldhitnnn = c.oldhitnnn;
	if (oldhitnnn2) *oldhitnnn2 = c.oldhitnnn2;
	if (hitnnn_last) *hitnnn_last = (int64_t)c.hitnnn_last;
	return CNF2_OK;
}

extern "C" int cnf2_negshift_minima(cnf2_engine* e, int32_t chrom, double* minval, int32_t* minstart)
{
	if (!e || !minval || !minstart) return CNF2_ERR_BAD_ARG;
	Slot* slp = nullptr;
	for (Slot& o : e->slots)
		if (chrom >= 0 && o.chrom == chrom) slp = &o;
	if (!slp) return e->fail(CNF2_ERR_STATE, "cnf2_negshift_minima: cnf2_estep for this chromosome first (its statistics must still be in flight)");
	Slot& sl = *slp;
	cudaSetDevice(e->settings.device);
	if (e->n_rows == 0) return CNF2_OK;
	CU(e->d_minval.ensure(e->n_rows));
	CU(e->d_minstart.ensure(e->n_rows));
	const int Mc = e->mc(chrom);
	negshift_minima_kernel<<<(unsigned)((e->n_rows * 32 + 127) / 128), 128, 0, sl.stream>>>(sl.stats_shared(), sl.stats_local(Mc), e->n_rows,
	                                                                                     e->n_shared, Mc, e->d_minval.p, e->d_minstart.p);
	e->n_kernels++;
	CU(cudaGetLastError());
	CU(cudaMemcpyAsync(minval, e->d_minval.p, (size_t)e->n_rows * sizeof(double), cudaMemcpyDeviceToHost, sl.stream));
	CU(cudaMemcpyAsync(minstart, e->d_minstart.p, (size_t)e->n_rows * sizeof(int32_t), cudaMemcpyDeviceToHost, sl.stream));
	CU(cudaStreamSynchronize(sl.stream));
	CU(cudaGetLastError());
	return CNF2_OK;
}

extern "C" int cnf2_apply_inversions(cnf2_engine* e, int32_t chrom, const int32_t* ids, const int32_t* minstart, int32_t n)
{
	if (!e || n < 0 || (n > 0 && (!ids || !minstart))) return CNF2_ERR_BAD_ARG;
	if (!e->have_ms_inputs) return e->fail(CNF2_ERR_STATE, "cnf2_apply_inversions: cnf2_set_mstep_inputs first");
	if (chrom < 0 || chrom + 1 >= (int)e->chromstarts.size()) return e->fail(CNF2_ERR_BAD_ARG, "cnf2_apply_inversions: bad chromosome");
	cudaSetDevice(e->settings.device);
	const int Mc = e->mc(chrom);
	std::vector<int32_t> rows, starts;
	for (int k = 0; k < n; k++)
	{
		if (ids[k] <= 0 || ids[k] >= e->n_ind) return e->fail(CNF2_ERR_BAD_ARG, "cnf2_apply_inversions: id out of range");
		if (minstart[k] < 0 || minstart[k] >= Mc) return e->fail(CNF2_ERR_BAD_ARG, "cnf2_apply_inversions: marker out of range");
		if (e->rowof[ids[k]] < 0) continue;  // not held by this rank
		rows.push_back(e->rowof[ids[k]]);
		starts.push_back(minstart[k]);
	}
	if (rows.empty()) return CNF2_OK;
	// the chromosome's haplobase / haplocount: the E-step's block while its pass has not run, the carried copy afterwards
	Slot* slp = nullptr;
	for (Slot& o : e->slots)
		if (o.chrom == chrom) slp = &o;
	const bool fresh = !e->carry_valid[chrom];
	if (fresh && !slp) return e->fail(CNF2_ERR_STATE, "cnf2_apply_inversions: no statistics of this chromosome (cnf2_estep first)");
	// behind the last pass / inversion (and, for fresh statistics, behind the chromosome's E-step); ahead of the next ones
	cudaStream_t st = e->stream;
	if (int rc = e->wait_chain(st)) return rc;
	if (slp)
	{
		CU(cudaEventRecord(slp->tail, slp->stream));
		CU(cudaStreamWaitEvent(st, slp->tail, 0));
	}
	CU(e->d_invrows.ensure(rows.size()));
	CU(e->d_invstart.ensure(rows.size()));
	CU(cudaMemcpyAsync(e->d_invrows.p, rows.data(), rows.size() * sizeof(int32_t), cudaMemcpyHostToDevice, st));
	CU(cudaMemcpyAsync(e->d_invstart.p, starts.data(), starts.size() * sizeof(int32_t), cudaMemcpyHostToDevice, st));
	MstepArgs a{};
	const size_t off = (size_t)e->n_rows * e->chromstarts[chrom];
	a.hw = e->d_hw.p + off;
	a.carry_base = e->d_carry_base.p + off;
	a.carry_count = e->d_carry_count.p + off;
	a.stats_shared = slp ? slp->stats_shared() : nullptr;
	a.stats_local = slp ? slp->stats_local(Mc) : nullptr;
	a.n_rows = e->n_rows;
	a.n_shared = e->n_shared;
	a.Mc = Mc;
	a.fresh = fresh ? 1 : 0;
	apply_inversions_kernel<<<(unsigned)rows.size(), 128, 0, st>>>(a, e->d_invrows.p, e->d_invstart.p, (int)rows.size());
	e->n_kernels++;
	CU(cudaGetLastError());
	if (int rc = e->mark_chain(st)) return rc;
	if (slp) CU(cudaStreamWaitEvent(slp->stream, e->ms_chain, 0));  // the chromosome's own pass reads what was just flipped
	CU(cudaStreamSynchronize(st));  // rows / starts are host vectors of this call
	return CNF2_OK;
}

// individ::arerelated (cnF2freq.cpp:916-946): within two generations up the pointer chains, or sharing an analysed child
struct RelStack
{
	int v[4], n = 0;
	bool has(int a) const
	{
		for (int i = 0; i < n; i++)
			if (v[i] == a) return true;
		return false;
	}
};
static bool arerelated(const cnf2_engine* e, int a, int b, RelStack stack, int gens)
{
	if (gens > 2) return false;
	if (b == a) return true;
	if (stack.has(a)) return false;
	stack.v[stack.n++] = a;
	if (stack.n == 1)
	{
		if (b && arerelated(e, b, a, stack, gens + 1)) return true;
	}
	for (int k = 0; k < 2; k++)
		if (const int p = e->parents[(size_t)a * 2 + k])
			if (arerelated(e, p, b, stack, gens + 1)) return true;
	for (int i : e->kids[a])
		for (int j : e->kids[b])
			if (i == j) return true;
	return false;
}

// Everybody `a` for whom arerelated(a, our) can hold -- a superset, to be confirmed by arerelated itself: our, its
// descendants and ancestors within two pointer steps, and whoever is tied to it through a shared analysed child at the levels
// arerelated looks at (the other parents of our's analysed children and of its parents' analysed children, and their
// descendants within two steps).  The reference scans all candidates of the chromosome for every individual; with
// thousands of candidates that scan is quadratic, this list is a handful of individuals except for a founder line.
static void related_superset(const cnf2_engine* e, int our, std::vector<int32_t>& out)
{
	out.clear();
	auto down2 = [&](int x) {
		out.push_back(x);
		for (int k1 : e->allkids[x])
		{
			out.push_back(k1);
			for (int k2 : e->allkids[k1]) out.push_back(k2);
		}
	};
	down2(our);
	for (int k = 0; k < 2; k++)
		if (const int p = e->parents[(size_t)our * 2 + k])
		{
			out.push_back(p);
			for (int kk = 0; kk < 2; kk++)
				if (const int pp = e->parents[(size_t)p * 2 + kk]) out.push_back(pp);
			for (int kid : e->kids[p])  // kids[q] and kids[a] share a child, q a parent of our
				for (int k2 = 0; k2 < 2; k2++)
					if (const int co = e->parents[(size_t)kid * 2 + k2]) out.push_back(co);
		}
	for (int kid : e->kids[our])  // kids[x] and kids[our] share a child for x = a, a parent of a, a grandparent of a
		for (int k2 = 0; k2 < 2; k2++)
			if (const int co = e->parents[(size_t)kid * 2 + k2]) down2(co);
	std::sort(out.begin(), out.end());
	out.erase(std::unique(out.begin(), out.end()), out.end());
}

extern "C" int cnf2_em_pass(cnf2_engine* e, int32_t chrom, int32_t* inv_ids, int32_t* inv_markers, int32_t cap, int32_t* n_inv)
{
	if (!e) return CNF2_ERR_BAD_ARG;
	if (!e->have_ms_inputs) return e->fail(CNF2_ERR_STATE, "cnf2_em_pass: cnf2_set_mstep_inputs first");
	const int C = (int)e->chromstarts.size() - 1;
	if (chrom < 0 || chrom >= C) return e->fail(CNF2_ERR_BAD_ARG, "cnf2_em_pass: bad chromosome");
	if ((int)e->candmap.size() != C) return e->fail(CNF2_ERR_STATE, "cnf2_em_pass: cnf2_begin_iteration first");
	Slot* slp = slot_of(e, chrom);
	if (!slp) return e->fail(CNF2_ERR_STATE, "cnf2_em_pass: cnf2_estep for this chromosome first (its statistics must still be in flight)");
	cudaSetDevice(e->settings.device);
	// minval / minstart of this chromosome for every individual (:6272-6284), the same on every rank
	{
		std::vector<double> mv(e->n_rows);
		std::vector<int32_t> ms(e->n_rows);
		if (int rc = cnf2_negshift_minima(e, chrom, mv.data(), ms.data())) return rc;
		const int Mc = e->mc(chrom);
		std::vector<double> g((size_t)e->n_ind * 2, 0.0);  // (minval, Mc - minstart): zero = no candidate
		const bool exchange = e->allreduce != nullptr;
		for (int r = exchange ? e->n_shared : 0; r < e->n_rows; r++)
			g[(size_t)e->rowid[r] * 2] = mv[r], g[(size_t)e->rowid[r] * 2 + 1] = (double)(Mc - ms[r]);
		if (exchange)
		{
			// rows only one rank holds: everybody else contributes zeros to the sum; shared rows are identical everywhere
			CU(e->d_gmin.ensure(g.size()));
			// on the control stream (see cnf2_mstep_pass): the host waits for this exchange, not for the E-steps issued ahead
			CU(cudaMemcpyAsync(e->d_gmin.p, g.data(), g.size() * sizeof(double), cudaMemcpyHostToDevice, e->comm2));
			if (e->allreduce(e->allreduce_ctx, e->d_gmin.p, g.size(), (void*)e->comm2)) return e->fail(CNF2_ERR_CUDA, "cnf2_em_pass: all-reduce hook failed");
			CU(cudaMemcpyAsync(g.data(), e->d_gmin.p, g.size() * sizeof(double), cudaMemcpyDeviceToHost, e->comm2));
			CU(cudaStreamSynchronize(e->comm2));
			for (int r = 0; r < e->n_shared; r++)
				g[(size_t)e->rowid[r] * 2] = mv[r], g[(size_t)e->rowid[r] * 2 + 1] = (double)(Mc - ms[r]);
		}
		e->minval_c[chrom].assign(e->n_ind, -1e-10);
		e->minstart_c[chrom].assign(e->n_ind, Mc);
		for (int id = 1; id < e->n_ind; id++)
			if (g[(size_t)id * 2 + 1] != 0)
			{
				e->minval_c[chrom][id] = g[(size_t)id * 2];
				e->minstart_c[chrom][id] = Mc - (int32_t)g[(size_t)id * 2 + 1];
				if (e->minstart_c[chrom][id] + 1 < Mc) e->hasmin[id] = 1;
			}
	}
	// candidates, at most one per individual and chromosome (:6270-6327): individuals in id order, every chromosome each
	std::vector<int32_t> rel;
	for (int id = 1; id < e->n_ind; id++)
	{
		if (!e->hasmin[id]) continue;  // no chromosome offers a candidate for this individual (:6301)
		bool have_rel = false;
		for (int c = 0; c < C; c++)
		{
			if (e->minval_c[c].empty()) continue;
			const int Mc = e->mc(c), minstart = e->minstart_c[c][id];
			if (!(minstart + 1 < Mc)) continue;  // :6301 (none, or the last marker)
			const double ourval = e->minval_c[c][id];
			auto& cm = e->candmap[c];
			bool anymatch = false;
			if (!cm.empty())
			{
				if (!have_rel) related_superset(e, id, rel), have_rel = true;
				// inferiorrelated (:3415-3435) over the candidates that can be related at all
				auto test = [&](std::map<int32_t, std::pair<double, int32_t>>::iterator it) {
					if (!arerelated(e, it->first, id, RelStack{}, 0)) return;
					if (it->second.first > ourval) cm.erase(it);
					else anymatch = true;
				};
				if (rel.size() < cm.size())
					for (int32_t a : rel)
					{
						auto it = cm.find(a);
						if (it != cm.end()) test(it);
					}
				else
					for (auto it = cm.begin(); it != cm.end();)
					{
						auto cur = it++;
						if (std::binary_search(rel.begin(), rel.end(), cur->first)) test(cur);
					}
			}
			if (!anymatch && rand() / (RAND_MAX / 5)) cm[id] = { ourval, minstart };  // :6319
		}
	}
	// negshifter on every candidate of every chromosome (:6336-6340)
	bool any = false;
	int n_out = 0;
	for (int c = 0; c < C; c++)
	{
		if (e->candmap[c].empty()) continue;
		any = true;
		std::vector<int32_t> ids, starts;
		for (const auto& cd : e->candmap[c])
		{
			ids.push_back(cd.first);
			starts.push_back(cd.second.second);
			if (inv_ids && inv_markers && n_out < cap) inv_ids[n_out] = cd.first, inv_markers[n_out] = (int32_t)e->chromstarts[c] + cd.second.second + 1;
			n_out++;
		}
		if (int rc = cnf2_apply_inversions(e, c, ids.data(), starts.data(), (int32_t)ids.size())) return rc;
	}
	if (n_inv) *n_inv = n_out;
	return cnf2_mstep_pass(e, chrom, any ? 1 : 0);
}

extern "C" int cnf2_get_params(cnf2_engine* e, int32_t chrom, uint8_t* alleles, double* sure, double* haploweight)
{
	if (!e) return CNF2_ERR_BAD_ARG;
	if (chrom < 0 || chrom >= (int)e->have_params.size() || !e->have_params[chrom])
		return e->fail(CNF2_ERR_STATE, "cnf2_get_params: no parameters of this chromosome on the device");
	cudaSetDevice(e->settings.device);
	const size_t M = e->markerpos.size(), c0 = e->chromstarts[chrom], Mc = e->mc(chrom);
	const size_t blk = (size_t)e->n_rows * (e->resident ? c0 : 0);
	if (int rc = e->wait_chain(e->stream)) return rc;
	for (const Run& r : e->runs)
	{
		const size_t dst = (size_t)r.id0 * M + c0, src = blk + (size_t)r.row0 * Mc;
		if (alleles)
			CU(cudaMemcpy2DAsync(alleles + dst * 2, M * 2, e->d_alleles.p + src, Mc * 2, Mc * 2, r.n, cudaMemcpyDeviceToHost, e->stream));
		if (sure) CU(cudaMemcpy2DAsync(sure + dst * 2, M * 16, e->d_sure.p + src, Mc * 16, Mc * 16, r.n, cudaMemcpyDeviceToHost, e->stream));
		if (haploweight)
			CU(cudaMemcpy2DAsync(haploweight + dst, M * 8, e->d_hw.p + src, Mc * 8, Mc * 8, r.n, cudaMemcpyDeviceToHost, e->stream));
	}
	CU(cudaStreamSynchronize(e->stream));
	return CNF2_OK;
}

extern "C" int cnf2_row_count(cnf2_engine* e, int32_t* n_rows)
{
	if (!e || !n_rows) return CNF2_ERR_BAD_ARG;
	*n_rows = e->n_rows;
	return CNF2_OK;
}

extern "C" int cnf2_row_ids(cnf2_engine* e, int32_t* ids)
{
	if (!e || !ids) return CNF2_ERR_BAD_ARG;
	for (int32_t r = 0; r < e->n_rows; r++) ids[r] = e->rowid[r];
	return CNF2_OK;
}

extern "C" int cnf2_get_stats_packed(cnf2_engine* e, int32_t chrom, double* haplobase, double* haplocount, double* infprobs,
                                     double* homozyg, double* loglik)
{
	if (!e) return CNF2_ERR_BAD_ARG;
	Slot* slp = nullptr;
	for (Slot& o : e->slots)
		if (chrom >= 0 && o.chrom == chrom) slp = &o;
	if (!slp) return e->fail(CNF2_ERR_STATE, "cnf2_get_stats_packed: cnf2_estep for this chromosome first (only the last E-steps in flight are kept)");
	Slot& sl = *slp;
	cudaSetDevice(e->settings.device);
	const size_t Mc = e->mc(chrom);
	// the device holds a shared block and a local block, each as planes haplobase | haplocount | infprobs | homozyg
	for (int blk = 0; blk < 2; blk++)
	{
		const size_t nr = blk ? (size_t)(e->n_rows - e->n_shared) : (size_t)e->n_shared, row0 = blk ? (size_t)e->n_shared : 0;
		if (!nr) continue;
		const double* base = blk ? sl.stats_local((int)Mc) : sl.stats_shared();
		const size_t plane = nr * Mc, off = row0 * Mc;
		if (haplobase) CU(cudaMemcpyAsync(haplobase + off, base, plane * 8, cudaMemcpyDeviceToHost, sl.stream));
		if (haplocount) CU(cudaMemcpyAsync(haplocount + off, base + plane, plane * 8, cudaMemcpyDeviceToHost, sl.stream));
		if (infprobs) CU(cudaMemcpyAsync(infprobs + off * 4, base + 2 * plane, plane * 32, cudaMemcpyDeviceToHost, sl.stream));
		if (homozyg) CU(cudaMemcpyAsync(homozyg + off * 2, base + 6 * plane, plane * 16, cudaMemcpyDeviceToHost, sl.stream));
	}
	if (loglik && !e->units.empty())
		CU(cudaMemcpyAsync(loglik, sl.loglik.p, e->units.size() * 2 * sizeof(double), cudaMemcpyDeviceToHost, sl.stream));
	CU(cudaStreamSynchronize(sl.stream));
	CU(cudaGetLastError());
	return CNF2_OK;
}

extern "C" int cnf2_get_negshift(cnf2_engine* e, int32_t chrom, double* negshift)
{
	if (!e || !negshift) return CNF2_ERR_BAD_ARG;
	Slot* slp = slot_of(e, chrom);
	if (!slp) return e->fail(CNF2_ERR_STATE, "cnf2_get_negshift: cnf2_estep for this chromosome first (only the E-steps in flight are kept)");
	Slot& sl = *slp;
	cudaSetDevice(e->settings.device);
	const size_t M = e->markerpos.size(), c0 = e->chromstarts[chrom], Mc = e->mc(chrom);
	for (const Run& r : e->runs)
	{
		const bool sh = r.row0 < e->n_shared;
		const StatsCell src = sh ? stats_cell(sl.stats_shared(), e->n_shared, Mc, r.row0, 0)
		                         : stats_cell(sl.stats_local((int)Mc), e->n_rows - e->n_shared, Mc, r.row0 - e->n_shared, 0);
		CU(cudaMemcpy2DAsync(negshift + (size_t)r.id0 * M + c0, M * 8, src.negshift, Mc * 8, Mc * 8, r.n, cudaMemcpyDeviceToHost, sl.stream));
	}
	CU(cudaStreamSynchronize(sl.stream));
	CU(cudaGetLastError());
	return CNF2_OK;
}

extern "C" int cnf2_get_negshift_packed(cnf2_engine* e, int32_t chrom, double* negshift)
{
	if (!e || !negshift) return CNF2_ERR_BAD_ARG;
	Slot* slp = slot_of(e, chrom);
	if (!slp) return e->fail(CNF2_ERR_STATE, "cnf2_get_negshift_packed: cnf2_estep for this chromosome first (only the E-steps in flight are kept)");
	Slot& sl = *slp;
	cudaSetDevice(e->settings.device);
	const size_t Mc = e->mc(chrom);
	for (int blk = 0; blk < 2; blk++)
	{
		const size_t nr = blk ? (size_t)(e->n_rows - e->n_shared) : (size_t)e->n_shared, row0 = blk ? (size_t)e->n_shared : 0;
		if (!nr) continue;
		const double* base = blk ? sl.stats_local((int)Mc) : sl.stats_shared();
		CU(cudaMemcpyAsync(negshift + row0 * Mc, base + 8 * nr * Mc, nr * Mc * 8, cudaMemcpyDeviceToHost, sl.stream));
	}
	CU(cudaStreamSynchronize(sl.stream));
	CU(cudaGetLastError());
	return CNF2_OK;
}

extern "C" int cnf2_estep_times(cnf2_engine* e, double* sum_ms, int64_t* n_launches, int64_t* n_kernels, int32_t reset)
{
	if (!e) return CNF2_ERR_BAD_ARG;
	cudaSetDevice(e->settings.device);
	for (Slot& sl : e->slots)
		if (int rc = e->collect_time(sl)) return rc;
	if (sum_ms) *sum_ms = e->estep_ms_sum;
	if (n_launches) *n_launches = e->estep_ms_n;
	if (n_kernels) *n_kernels = e->n_kernels;
	if (reset) e->estep_ms_sum = 0, e->estep_ms_n = 0, e->n_kernels = 0;
	return CNF2_OK;
}

extern "C" int cnf2_mstep_counters(cnf2_engine* e, uint64_t* evals, int32_t reset)
{
	if (!e) return CNF2_ERR_BAD_ARG;
	if (!e->have_ms_inputs) return e->fail(CNF2_ERR_STATE, "cnf2_mstep_counters: cnf2_set_mstep_inputs first");
	cudaSetDevice(e->settings.device);
	if (int rc = cnf2_sync(e)) return rc;
	EmCtl c{};
	CU(cudaMemcpy(&c, e->d_ctl.p, sizeof(c), cudaMemcpyDeviceToHost));
	if (evals) evals[0] = c.evals[0], evals[1] = c.evals[1], evals[2] = c.closing[0], evals[3] = c.closing[1];
	if (reset) CU(cudaMemset(&e->d_ctl.p->evals[0], 0, sizeof(c.evals) + sizeof(c.closing)));
	return CNF2_OK;
}

extern "C" int cnf2_get_params_packed(cnf2_engine* e, int32_t chrom, uint8_t* alleles, double* sure, double* haploweight)
{
	if (!e) return CNF2_ERR_BAD_ARG;
	if (chrom < 0 || chrom >= (int)e->have_params.size() || !e->have_params[chrom])
		return e->fail(CNF2_ERR_STATE, "cnf2_get_params_packed: no parameters of this chromosome on the device");
	cudaSetDevice(e->settings.device);
	const size_t c0 = e->chromstarts[chrom], Mc = e->mc(chrom), n = (size_t)e->n_rows * Mc;
	const size_t blk = (size_t)e->n_rows * (e->resident ? c0 : 0);
	if (int rc = e->wait_chain(e->stream)) return rc;
	if (alleles) CU(cudaMemcpyAsync(alleles, e->d_alleles.p + blk, n * 2, cudaMemcpyDeviceToHost, e->stream));
	if (sure) CU(cudaMemcpyAsync(sure, e->d_sure.p + blk, n * 16, cudaMemcpyDeviceToHost, e->stream));
	if (haploweight) CU(cudaMemcpyAsync(haploweight, e->d_hw.p + blk, n * 8, cudaMemcpyDeviceToHost, e->stream));
	CU(cudaStreamSynchronize(e->stream));
	return CNF2_OK;
}

extern "C" int cnf2_get_params_packed_async(cnf2_engine* e, int32_t chrom, uint8_t* alleles, double* sure, double* haploweight)
{
	if (!e) return CNF2_ERR_BAD_ARG;
	if (chrom < 0 || chrom >= (int)e->have_params.size() || !e->have_params[chrom])
		return e->fail(CNF2_ERR_STATE, "cnf2_get_params_packed_async: no parameters of this chromosome on the device");
	cudaSetDevice(e->settings.device);
	const size_t c0 = e->chromstarts[chrom], Mc = e->mc(chrom), n = (size_t)e->n_rows * Mc;
	const size_t blk = (size_t)e->n_rows * (e->resident ? c0 : 0);
	// behind the last pass / inversion enqueued so far (what it wrote is what is read), beside whatever is enqueued later
	if (int rc = e->wait_chain(e->copy)) return rc;
	if (alleles) CU(cudaMemcpyAsync(alleles, e->d_alleles.p + blk, n * 2, cudaMemcpyDeviceToHost, e->copy));
	if (sure) CU(cudaMemcpyAsync(sure, e->d_sure.p + blk, n * 16, cudaMemcpyDeviceToHost, e->copy));
	if (haploweight) CU(cudaMemcpyAsync(haploweight, e->d_hw.p + blk, n * 8, cudaMemcpyDeviceToHost, e->copy));
	return CNF2_OK;
}

extern "C" int cnf2_last_estep_ms(cnf2_engine* e, float* ms)
{
	if (!e || !ms || e->last_slot < 0) return CNF2_ERR_BAD_ARG;
	cudaSetDevice(e->settings.device);
	Slot& sl = e->slots[e->last_slot];
	CU(cudaEventSynchronize(sl.ev1));
	CU(cudaEventElapsedTime(ms, sl.ev0, sl.ev1));
	return CNF2_OK;
}

extern "C" int cnf2_last_estep_launches(cnf2_engine* e, int32_t* n)
{
	if (!e || !n) return CNF2_ERR_BAD_ARG;
	*n = e->last_launches;
	return CNF2_OK;
}

extern "C" int cnf2_mark(cnf2_engine* e, int32_t slot)
{
	if (!e || slot < 0 || slot >= 4) return CNF2_ERR_BAD_ARG;
	cudaSetDevice(e->settings.device);
	if (!e->marks[slot]) CU(cudaEventCreate(&e->marks[slot]));
	// the mark comes after everything enqueued so far and before everything enqueued later, on every stream
	for (Slot& sl : e->slots)
	{
		CU(cudaEventRecord(sl.tail, sl.stream));
		CU(cudaStreamWaitEvent(e->stream, sl.tail, 0));
	}
	CU(cudaEventRecord(e->marks[slot], e->stream));
	for (Slot& sl : e->slots) CU(cudaStreamWaitEvent(sl.stream, e->marks[slot], 0));
	return CNF2_OK;
}

extern "C" int cnf2_elapsed_ms(cnf2_engine* e, int32_t from, int32_t to, float* ms)
{
	if (!e || !ms || from < 0 || from >= 4 || to < 0 || to >= 4 || !e->marks[from] || !e->marks[to]) return CNF2_ERR_BAD_ARG;
	cudaSetDevice(e->settings.device);
	CU(cudaEventSynchronize(e->marks[to]));
	CU(cudaEventElapsedTime(ms, e->marks[from], e->marks[to]));
	return CNF2_OK;
}

extern "C" void* cnf2_control_stream(cnf2_engine* e)
{
	return e ? (void*)e->comm2 : nullptr;
}

extern "C" void* cnf2_stream(cnf2_engine* e)
{
	return e ? (void*)e->stream : nullptr;
}

// ---------------------------------------------------------------------------------------------------
// Roofline denominators measured on the box itself: fp64 FMA issue rate and a device copy.
namespace
{
__global__ void dfma_peak_kernel(double* out, int iters, double a, double b)
{
	double x0 = threadIdx.x, x1 = x0 + 1, x2 = x0 + 2, x3 = x0 + 3, x4 = x0 + 4, x5 = x0 + 5, x6 = x0 + 6, x7 = x0 + 7;
	for (int i = 0; i < iters; i++)
	{
		x0 = fma(x0, a, b);
		x1 = fma(x1, a, b);
		x2 = fma(x2, a, b);
		x3 = fma(x3, a, b);
		x4 = fma(x4, a, b);
		x5 = fma(x5, a, b);
		x6 = fma(x6, a, b);
		x7 = fma(x7, a, b);
	}
	out[blockIdx.x * blockDim.x + threadIdx.x] = x0 + x1 + x2 + x3 + x4 + x5 + x6 + x7;
}

__global__ void copy_peak_kernel(const double2* __restrict__ src, double2* __restrict__ dst, size_t n)
{
	for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) dst[i] = src[i];
}
}  // namespace

extern "C" int cnf2_measure_peaks(cnf2_engine* e, double* dfma_tflops, double* copy_gbs)
{
	if (!e) return CNF2_ERR_BAD_ARG;
	cudaSetDevice(e->settings.device);
	cudaEvent_t t0, t1;
	CU(cudaEventCreate(&t0));
	CU(cudaEventCreate(&t1));
	float ms = 0;
	if (dfma_tflops)
	{
		const int blocks = e->n_sm * 8, threads = 256, iters = 1 << 14;
		double* out = nullptr;
		CU(cudaMalloc(&out, (size_t)blocks * threads * sizeof(double)));
		double best = 0;
		for (int rep = 0; rep < 4; rep++)
		{
			CU(cudaEventRecord(t0, e->stream));
			dfma_peak_kernel<<<blocks, threads, 0, e->stream>>>(out, iters, 0.999999, 1e-9);
			CU(cudaEventRecord(t1, e->stream));
			CU(cudaEventSynchronize(t1));
			CU(cudaEventElapsedTime(&ms, t0, t1));
			const double flops = 2.0 * 8 * (double)iters * blocks * threads;
			if (rep) best = std::max(best, flops / (ms * 1e-3) * 1e-12);
		}
		cudaFree(out);
		*dfma_tflops = best;
	}
	if (copy_gbs)
	{
		const size_t n = (size_t)1 << 26;  // 2 x 1 GiB
		double2 *a = nullptr, *b = nullptr;
		CU(cudaMalloc(&a, n * sizeof(double2)));
		CU(cudaMalloc(&b, n * sizeof(double2)));
		CU(cudaMemsetAsync(a, 0, n * sizeof(double2), e->stream));
		double best = 0;
		for (int rep = 0; rep < 5; rep++)
		{
			CU(cudaEventRecord(t0, e->stream));
			copy_peak_kernel<<<e->n_sm * 16, 512, 0, e->stream>>>(a, b, n);
			CU(cudaEventRecord(t1, e->stream));
			CU(cudaEventSynchronize(t1));
			CU(cudaEventElapsedTime(&ms, t0, t1));
			if (rep) best = std::max(best, 2.0 * n * sizeof(double2) / (ms * 1e-3) * 1e-9);
		}
		cudaFree(a);
		cudaFree(b);
		*copy_gbs = best;
	}
	cudaEventDestroy(t0);
	cudaEventDestroy(t1);
	return CNF2_OK;
}

// ---------------------------------------------------------------------------------------------------
// Row f3: postmarkerdata (cnF2freq.cpp:3191-3412).  The emission scans run on the device (cnf2_preprocess.cuh); the
// sequential bookkeeping between them -- fixkid, the descendant counts, the per-marker allele evidence fixparents hands
// to the parents and the corrections drawn from it, lockhaplos -- follows the reference's single-threaded order
// (individuals by id, markers inside) on the host.
#include "cnf2_preprocess.cuh"

namespace
{
// individ::markervals[marker]: allele -> (count, probability), ordered by allele value like the reference's map; at most a
// handful of distinct alleles ever appear
struct AlleleEvidence
{
	struct Entry
	{
		int key, count;
		double prob;
	};
	Entry e[4];
	int n = 0;
	Entry* find(int key)
	{
		for (int i = 0; i < n; i++)
			if (e[i].key == key) return &e[i];
		return nullptr;
	}
	bool insert(int key, int count, double prob)  // map::insert: keeps an existing entry
	{
		if (find(key)) return true;
		if (n >= 4) return false;
		int at = n++;
		while (at > 0 && e[at - 1].key > key) e[at] = e[at - 1], at--;
		e[at] = Entry{ key, count, prob };
		return true;
	}
	void erase(int key)
	{
		for (int i = 0; i < n; i++)
			if (e[i].key == key)
			{
				for (int j = i + 1; j < n; j++) e[j - 1] = e[j];
				n--;
				return;
			}
	}
};

double dosureval(int what, int count, double prob)  // :3082-3097
{
	(void)count;
	if (prob == 0) return 0;
	double toret = std::log(prob);
	toret /= what;
	toret *= 4;
	toret = std::exp(toret);
	toret /= (1 + toret);
	return toret;
}
}  // namespace

extern "C" int cnf2_preprocess(cnf2_engine* e, int32_t n_ind, const int32_t* parents, const uint8_t* exists, const uint8_t* empty,
                               uint8_t* alleles, double* sure, double* haploweight, int32_t* descendants, int32_t* children,
                               uint8_t* founder, cnf2_log_fn log, void* log_ctx)
{
	if (!e || !parents || !exists || !empty || !alleles || !sure || !haploweight || !descendants || !children || !founder || n_ind <= 0)
		return e ? e->fail(CNF2_ERR_BAD_ARG, "cnf2_preprocess: null or empty argument") : CNF2_ERR_BAD_ARG;
	if (e->chromstarts.empty()) return e->fail(CNF2_ERR_STATE, "cnf2_preprocess: cnf2_set_map first");
	cudaSetDevice(e->settings.device);
	const size_t K = n_ind, M = e->markerpos.size(), cells = K * M;
	auto say = [&](int stream, const char* fmt, auto... args) {
		if (!log) return;
		char line[256];
		snprintf(line, sizeof(line), fmt, args...);
		log(log_ctx, stream, line);
	};
	DevBuf<uchar2> d_al;
	DevBuf<double2> d_su;
	DevBuf<double> d_hw, d_sq, d_sum;
	DevBuf<int2> d_par;
	DevBuf<uint8_t> d_exists, d_steady, d_ok, d_ign;
	auto cleanup = [&]() {
		d_al.release(), d_su.release(), d_hw.release(), d_sq.release(), d_sum.release(), d_par.release();
		d_exists.release(), d_steady.release(), d_ok.release(), d_ign.release();
	};
#define PRE(call)                                    \
	do                                               \
	{                                                \
		int rc_ = e->cuda((call), #call);            \
		if (rc_)                                     \
		{                                            \
			cleanup();                               \
			return rc_;                              \
		}                                            \
	} while (0)
	PRE(d_al.ensure(cells));
	PRE(d_su.ensure(cells));
	PRE(d_hw.ensure(cells));
	PRE(d_par.ensure(K));
	PRE(d_exists.ensure(K));
	PRE(d_steady.ensure(K));
	PRE(d_ok.ensure(cells));
	PRE(d_ign.ensure(K));
	PRE(cudaMemcpy(d_par.p, parents, K * sizeof(int2), cudaMemcpyHostToDevice));
	PRE(cudaMemcpy(d_exists.p, exists, K, cudaMemcpyHostToDevice));
	// what fixtrees leaves in individ::founder (:3170-3177) and returns as flag2ignore (:3105-3167) for ANY individual
	std::vector<uint8_t> steady(K, 0), ign(K, 0);
	for (size_t ind = 1; ind < K; ind++)
	{
		if (!exists[ind]) continue;
		int flag2ignore = 1;
		bool any1 = false;
		for (int lev1 = 0; lev1 < 2; lev1++)
		{
			const int l1 = parents[ind * 2 + lev1];
			if (!l1) continue;
			const int flag2index = 1 + lev1 * 3;
			if (!empty[l1]) flag2ignore |= 1 << flag2index;
			bool anypars = false;
			for (int lev2 = 0; lev2 < 2; lev2++)
			{
				const int l2 = parents[(size_t)l1 * 2 + lev2];
				if (!l2 || empty[l2]) continue;
				flag2ignore |= 1 << (flag2index + lev2 + 1);
				anypars = true;
			}
			if (anypars || !empty[l1]) any1 = true;
		}
		steady[ind] = any1 ? 0 : 1;
		ign[ind] = (uint8_t)(flag2ignore ^ (NUMPATHS - 1));
	}
	PRE(cudaMemcpy(d_steady.p, steady.data(), K, cudaMemcpyHostToDevice));
	PRE(cudaMemcpy(d_ign.p, ign.data(), K, cudaMemcpyHostToDevice));
	PreArgs pa{};
	pa.alleles = d_al.p, pa.sure = d_su.p, pa.hw = d_hw.p, pa.par = d_par.p, pa.steady = d_steady.p;
	pa.K = (int32_t)K, pa.M = (int32_t)M;
	auto upload = [&]() -> cudaError_t {
		cudaError_t ce = cudaMemcpy(d_al.p, alleles, cells * 2, cudaMemcpyHostToDevice);
		if (ce == cudaSuccess) ce = cudaMemcpy(d_su.p, sure, cells * 16, cudaMemcpyHostToDevice);
		if (ce == cudaSuccess) ce = cudaMemcpy(d_hw.p, haploweight, cells * 8, cudaMemcpyHostToDevice);
		return ce;
	};
	auto md = [&](size_t ind, size_t g, int side) -> uint8_t& { return alleles[(ind * M + g) * 2 + side]; };
	auto ms = [&](size_t ind, size_t g, int side) -> double& { return sure[(ind * M + g) * 2 + side]; };
	std::vector<AlleleEvidence> vals(cells);
	std::vector<uint8_t> okmask(cells);
	int any = 0, anyrem = 0;
	do
	{
		for (size_t i = 1; i < K; i++) children[i] = 0;
		// fixkid (:1469-1487): an unknown genotype takes the allele of a homozygous parent, sureness 0.5
		for (size_t i = 1; i < K; i++)
		{
			if (!exists[i]) continue;
			for (size_t g = 0; g < M; g++)
			{
				if (md(i, g, 0) != UNKNOWN || md(i, g, 1) != UNKNOWN) continue;
				for (int p = 0; p < 2; p++)
				{
					const int par = parents[i * 2 + p];
					if (!par || !exists[par]) continue;
					if (md(par, g, 0) == UNKNOWN) continue;
					if (md(par, g, 0) != md(par, g, 1)) continue;
					md(i, g, p) = md(par, g, 0);
					ms(i, g, p) = 0.5;
				}
			}
		}
		// descendants (:3228-3258): every individual sends its count (at least 1) up once more per round
		{
			std::vector<int> upsent(K, 0);
			bool anychange;
			do
			{
				anychange = false;
				for (size_t i = 1; i < K; i++)
				{
					if (!exists[i]) continue;
					int nowdesc = descendants[i];
					if (!nowdesc) nowdesc = 1;
					nowdesc -= upsent[i];
					if (nowdesc > 0)
					{
						for (int p = 0; p < 2; p++)
							if (const int par = parents[i * 2 + p])
								if (exists[par]) descendants[par] += nowdesc;
						upsent[i] += nowdesc;
						anychange = true;
					}
				}
			} while (anychange);
			for (size_t i = 1; i < K; i++)
				if (exists[i] && descendants[i] == 0) descendants[i] = 1;
		}
		// fixparents' scan on the device.  A genotype without any admissible interpretation is cleared (:1432-1437), which the
		// scans of its descendants with a higher id see: the pass is repeated from the first clearing of a marker onwards.
		std::vector<size_t> final_upto(M, 1);
		std::vector<uint8_t> cleared(cells, 0);
		for (;;)
		{
			PRE(upload());
			if (e->selfing) pre_okmask_kernel<true><<<(unsigned)((cells + 127) / 128), 128, 0, e->stream>>>(pa, d_exists.p, d_ok.p);
			else pre_okmask_kernel<false><<<(unsigned)((cells + 127) / 128), 128, 0, e->stream>>>(pa, d_exists.p, d_ok.p);
			e->n_kernels++;
			PRE(cudaGetLastError());
			PRE(cudaMemcpyAsync(okmask.data(), d_ok.p, cells, cudaMemcpyDeviceToHost, e->stream));
			PRE(cudaStreamSynchronize(e->stream));
			bool again = false;
			for (size_t g = 0; g < M; g++)
			{
				size_t i = final_upto[g];
				for (; i < K; i++)
					if (exists[i] && !okmask[i * M + g] && !cleared[i * M + g]) break;
				if (i < K)
				{
					say(1, "Clearing %d:%d (was %d,%d)\n", (int)i, (int)g, (int)md(i, g, 0), (int)md(i, g, 1));
					md(i, g, 0) = md(i, g, 1) = UNKNOWN;
					ms(i, g, 0) = ms(i, g, 1) = 0.0;
					cleared[i * M + g] = 1;
					final_upto[g] = i + 1;
					again = true;
				}
				else final_upto[g] = K;
			}
			if (!again) break;
		}
		// the evidence fixparents hands to the parents (:1439-1466), individuals in id order
		for (size_t i = 1; i < K; i++)
		{
			if (!exists[i]) continue;
			for (int j = 0; j < 2; j++)
				if (const int par = parents[i * 2 + j])
					if (exists[par]) children[par]++;
			for (size_t g = 0; g < M; g++)
			{
				const bool ok[2] = { (okmask[i * M + g] & 1) != 0 && !cleared[i * M + g], (okmask[i * M + g] & 2) != 0 && !cleared[i * M + g] };
				if (ok[0] == ok[1]) continue;
				for (int flag2 = 0; flag2 < 2; flag2++)
				{
					if (!ok[flag2]) continue;
					for (int k = 0; k < 2; k++)
					{
						const int par = parents[i * 2 + k];
						if (!par || !exists[par]) continue;
						const int u = (k ^ flag2) & 1;
						const int allele = md(i, g, u);
						if (allele == UNKNOWN) continue;
						double old1 = 1;
						int old2 = 0;
						AlleleEvidence& ev = vals[(size_t)par * M + g];
						if (AlleleEvidence::Entry* en = ev.find(allele)) old1 = en->prob, old2 = en->count;
						double probit = ms(i, g, 0) + ms(i, g, 1);
						probit /= (1 - probit);
						if (AlleleEvidence::Entry* en = ev.find(allele)) en->count = old2 + 1, en->prob = old1 * probit;
						else ev.insert(allele, old2 + 1, old1 * probit);
					}
				}
			}
		}
		// corrections (:3283-3351)
		any = 0;
		anyrem = 0;
		for (size_t i = 1; i < K; i++)
		{
			if (!exists[i]) continue;
			for (size_t g = 0; g < M; g++)
			{
				AlleleEvidence& ev = vals[i * M + g];
				const int known = (md(i, g, 0) != UNKNOWN) + (md(i, g, 1) != UNKNOWN);
				const int oldany = any;
				if (known == 2) continue;
				ev.erase(UNKNOWN);
				if (md(i, g, 0) != UNKNOWN) ev.insert(md(i, g, 0), children[i], ms(i, g, 0));
				if (md(i, g, 1) != UNKNOWN) ev.insert(md(i, g, 1), children[i], ms(i, g, 1));
				if (ev.n >= 3) say(1, "Error, too many matches: %d\t%d\n", (int)i, (int)g);
				if (ev.n == 2)
				{
					const int knowncount = ev.e[0].count + ev.e[1].count;
					md(i, g, 0) = (uint8_t)ev.e[0].key;
					md(i, g, 1) = (uint8_t)ev.e[1].key;
					ms(i, g, 0) = dosureval(knowncount, ev.e[0].count, ev.e[0].prob);
					ms(i, g, 1) = dosureval(knowncount, ev.e[1].count, ev.e[1].prob);
					any++;
				}
				else if (ev.n == 1 && known == 0)
				{
					any++;
					md(i, g, 0) = (uint8_t)ev.e[0].key;
					md(i, g, 1) = UNKNOWN;
					ms(i, g, 0) = dosureval(ev.e[0].count, ev.e[0].count, ev.e[0].prob);
					ms(i, g, 1) = 0.0;
				}
				if (any != oldany)
					say(0, "Correction at %d, marker %d (%d;%d) (%lf;%lf)\n", (int)i, (int)g, (int)md(i, g, 0), (int)md(i, g, 1), ms(i, g, 0), ms(i, g, 1));
			}
			for (size_t g = 0; g < M; g++)
			{
				if (md(i, g, 0) == SEXMARKER) std::swap(md(i, g, 0), md(i, g, 1));
				vals[i * M + g].n = 0;
			}
		}
		say(1, "Number of corrected genotypes: %d\n", any);
	} while (any > anyrem);

	// addvariance (:1489-1558) for every individual (fixtrees' sticky founder flag takes effect in id order, :3377), then
	// lockhaplos (:3045-3079): the phase is locked at the marker of largest variance of each chromosome
	for (size_t i = 1; i < K; i++) founder[i] = exists[i] ? steady[i] : 0;
	PRE(upload());
	const size_t C = e->chromstarts.size() - 1;
	// Which cells the literal sums are needed for.  SELFING: all of them (the factorised form is written for 64 states).
	// Otherwise pre_variance_fast_kernel ranks the markers of every (individual, chromosome); only a marker within 1e-9 of the
	// largest value can hold lockhaplos' strict maximum (the two forms differ by rounding, ~1e-15), and among markers with the
	// same genotypes and sureness in the whole focus tree (equal signatures: bit-identical literal sums) only the first can.
	// CNF2_PRE_LITERAL=1 forces the literal kernel everywhere (A/B switch of the tests).
	const char* force_lit = getenv("CNF2_PRE_LITERAL");
	const bool all_literal = e->selfing || (force_lit && force_lit[0] == '1');
	std::vector<unsigned long long> list;
	if (!all_literal)
	{
		DevBuf<double> d_rank;
		DevBuf<unsigned long long> d_sig;
		auto drop = [&]() { d_rank.release(), d_sig.release(); };
		int rc = e->cuda(d_rank.ensure(cells), "cudaMalloc(rank)");
		if (!rc) rc = e->cuda(d_sig.ensure(cells), "cudaMalloc(signatures)");
		if (!rc)
		{
			pre_variance_fast_kernel<<<(unsigned)((cells + 127) / 128), 128, 0, e->stream>>>(pa, d_exists.p, d_ign.p, d_rank.p, d_sig.p);
			e->n_kernels++;
			rc = e->cuda(cudaGetLastError(), "pre_variance_fast_kernel");
		}
		std::vector<double> rank(cells);
		std::vector<unsigned long long> sig(cells);
		if (!rc) rc = e->cuda(cudaMemcpyAsync(rank.data(), d_rank.p, cells * sizeof(double), cudaMemcpyDeviceToHost, e->stream), "cudaMemcpy(rank)");
		if (!rc) rc = e->cuda(cudaMemcpyAsync(sig.data(), d_sig.p, cells * sizeof(unsigned long long), cudaMemcpyDeviceToHost, e->stream), "cudaMemcpy(signatures)");
		if (!rc) rc = e->cuda(cudaStreamSynchronize(e->stream), "cudaStreamSynchronize");
		drop();
		if (rc)
		{
			cleanup();
			return rc;
		}
		std::vector<unsigned long long> seen;
		for (size_t i = 1; i < K; i++)
		{
			if (!exists[i]) continue;
			for (size_t c = 0; c < C; c++)
			{
				const size_t j0 = e->chromstarts[c], j1 = e->chromstarts[c + 1];
				double best = 0;
				bool noisy = false;
				for (size_t j = j0; j < j1; j++)
				{
					const double v = std::fabs(rank[i * M + j]);
					if (v > best) best = v, noisy = rank[i * M + j] < 0;
				}
				if (!(best > 0)) continue;  // every marker's two alleles give the same sums: no variance anywhere, nothing is locked
				// a largest value at rounding-noise level ranks nothing: every marker with any variance is then evaluated literally
				const double thr = noisy ? 0.0 : best * (1.0 - 1e-9);
				seen.clear();
				for (size_t j = j0; j < j1; j++)
				{
					const double v = std::fabs(rank[i * M + j]);
					if (!(v > 0) || v < thr) continue;
					const unsigned long long h = sig[i * M + j];
					if (std::find(seen.begin(), seen.end(), h) != seen.end()) continue;
					seen.push_back(h);
					list.push_back((unsigned long long)(i * M + j));
				}
			}
		}
	}
	const size_t n_eval = all_literal ? cells : list.size();
	DevBuf<unsigned long long> d_list;
	std::vector<double> psq(n_eval * 8), psum(n_eval * 8);
	if (n_eval > 0)
	{
		PRE(d_sq.ensure(n_eval * 8));
		PRE(d_sum.ensure(n_eval * 8));
		if (!all_literal)
		{
			int rc = e->cuda(d_list.ensure(n_eval), "cudaMalloc(list)");
			if (!rc) rc = e->cuda(cudaMemcpyAsync(d_list.p, list.data(), n_eval * sizeof(unsigned long long), cudaMemcpyHostToDevice, e->stream), "cudaMemcpy(list)");
			if (rc)
			{
				d_list.release();
				cleanup();
				return rc;
			}
		}
		const unsigned grid = (unsigned)((n_eval * 8 + 127) / 128);
		if (e->selfing) pre_variance_kernel<true><<<grid, 128, 0, e->stream>>>(pa, d_exists.p, d_ign.p, d_sq.p, d_sum.p, nullptr, 0);
		else pre_variance_kernel<false><<<grid, 128, 0, e->stream>>>(pa, d_exists.p, d_ign.p, d_sq.p, d_sum.p, all_literal ? nullptr : d_list.p, n_eval);
		e->n_kernels++;
		int rc = e->cuda(cudaGetLastError(), "pre_variance_kernel");
		if (!rc) rc = e->cuda(cudaMemcpyAsync(psq.data(), d_sq.p, n_eval * 8 * sizeof(double), cudaMemcpyDeviceToHost, e->stream), "cudaMemcpy(sq)");
		if (!rc) rc = e->cuda(cudaMemcpyAsync(psum.data(), d_sum.p, n_eval * 8 * sizeof(double), cudaMemcpyDeviceToHost, e->stream), "cudaMemcpy(sum)");
		if (!rc) rc = e->cuda(cudaStreamSynchronize(e->stream), "cudaStreamSynchronize");
		d_list.release();
		if (rc)
		{
			cleanup();
			return rc;
		}
	}
	e->pre_literal_cells = n_eval;
	// lockhaplos (:3045-3079): the first marker holding the strict maximum of the literal variances
	size_t at = 0;  // position in list (ascending in (individual, marker))
	for (size_t i = 1; i < K; i++)
	{
		if (!exists[i]) continue;
		for (size_t c = 0; c < C; c++)
		{
			long bestpos = -1;
			double bestvar = 0;
			auto consider = [&](size_t j, size_t slot) {
				double sum = 0, sqsum = 0;
				for (int k = 0; k < 8; k++) sum += psum[slot * 8 + k], sqsum += psq[slot * 8 + k];
				const double variance = sum ? sqsum : 0.0;  // `if (!sum) return` leaves the vector's zero (:1553)
				if (variance > bestvar) bestpos = (long)j, bestvar = variance;
			};
			if (all_literal)
				for (size_t j = e->chromstarts[c]; j != e->chromstarts[c + 1]; j++) consider(j, i * M + j);
			else
				for (; at < list.size() && list[at] < (unsigned long long)(i * M + e->chromstarts[c + 1]); at++) consider((size_t)(list[at] - i * M), at);
			if (bestpos == -1) continue;
			say(0, "Fixing point: %d %d %d\n", (int)i, (int)c + 1, (int)bestpos);
			haploweight[i * M + bestpos] = haploweight[i * M + bestpos] <= 0.5 ? 0 : 1;
		}
	}
	cleanup();
#undef PRE
	return CNF2_OK;
}

extern "C" int64_t cnf2_preprocess_literal_cells(const cnf2_engine* e)
{
	return e ? (int64_t)e->pre_literal_cells : -1;
}

// ---- an entry point that allocates nothing (no counterpart in the reference) ----
// cnf2_negshift_minima_noalloc: the minima of cnf2_negshift_minima for callers that cannot call cudaMalloc (CUDA graph
// capture, serving loops, harnesses that forbid it).  Inputs and outputs are the caller's device buffers; the kernels'
// scratch -- one partial minimum per (row, chunk of NA_CHUNK markers) -- is the file-scope array below, held by every GPU
// that loads the library until the process exits.  Two kernels: chunk minima into the scratch (several warps share a long
// row instead of one warp walking it alone), then per row the minimum of its chunks and the first marker holding it, found
// by skipping the chunks whose partial minimum rules them out.  Every scratch cell that the second kernel reads is
// overwritten by the first, so nothing has to start at zero and the previous call's contents do not matter; all calls share
// the array, so calls that overlap on the GPU (different streams, no dependency) race on it.
constexpr int NA_CHUNK = 256;                 // markers per scratch cell
constexpr size_t NA_MAX_CELLS = 1u << 20;     // 8 MiB of scratch: n_rows * ceil(Mc / NA_CHUNK) may not exceed this
static __device__ __align__(256) double na_chunkmin[NA_MAX_CELLS];
namespace
{

__global__ void __launch_bounds__(128) na_chunk_minima_kernel(const double* __restrict__ ns, int n_rows, int Mc, int nchunk, double* __restrict__ chunkmin)
{
	const size_t cell = ((size_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
	const int lane = threadIdx.x & 31;
	if (cell >= (size_t)n_rows * nchunk) return;
	const int row = (int)(cell / nchunk), ch = (int)(cell % nchunk);
	const double* r = ns + (size_t)row * Mc;
	const int mb = ch * NA_CHUNK, me = min(Mc, mb + NA_CHUNK);
	double best = -1e-10;
	for (int m = mb + lane; m < me; m += 32) best = fmin(best, r[m]);
	for (int o = 16; o > 0; o >>= 1) best = fmin(best, __shfl_xor_sync(0xffffffffu, best, o));
	if (lane == 0) chunkmin[cell] = best;
}

// same decisions as negshift_minima_kernel: the minimum below -1e-10, the first marker within the scores' rounding noise of it
__global__ void __launch_bounds__(128) na_row_minima_kernel(const double* __restrict__ ns, int n_rows, int Mc, int nchunk, const double* __restrict__ chunkmin,
                                                            double* __restrict__ minval, int32_t* __restrict__ minstart)
{
	const int row = (int)(((size_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5), lane = threadIdx.x & 31;
	if (row >= n_rows) return;
	const double* cm = chunkmin + (size_t)row * nchunk;
	double best = -1e-10;
	for (int c = lane; c < nchunk; c += 32) best = fmin(best, cm[c]);
	for (int o = 16; o > 0; o >>= 1) best = fmin(best, __shfl_xor_sync(0xffffffffu, best, o));
	int at = Mc;
	double val = best;
	if (best < -1e-10)
	{
		const double thr = best + fmax(1e-15, -best * 1e-12);
		int first = nchunk;  // the first chunk with a score at or below the threshold (one exists: the minimum's)
		for (int c0 = 0; c0 < nchunk && first == nchunk; c0 += 32)
		{
			const int c = c0 + lane;
			const unsigned hit = __ballot_sync(0xffffffffu, c < nchunk && cm[c] <= thr);
			if (hit) first = c0 + __ffs(hit) - 1;
		}
		const double* r = ns + (size_t)row * Mc;
		const int me = min(Mc, (first + 1) * NA_CHUNK);
		for (int m0 = first * NA_CHUNK; m0 < me && at == Mc; m0 += 32)
		{
			const int m = m0 + lane;
			const double v = m < me ? r[m] : 0.0;
			const unsigned hit = __ballot_sync(0xffffffffu, m < me && v <= thr);
			if (hit)
			{
				const int src = __ffs(hit) - 1;
				at = m0 + src;
				val = __shfl_sync(0xffffffffu, v, src);
			}
		}
	}
	if (lane == 0) minval[row] = val, minstart[row] = at;
}

// the scratch array's device address, looked up once per device with that device current (in host code the array's name is
// only a stand-in for the symbol, not an address a kernel could use)
constexpr int NA_MAX_DEVICES = 64;
double* na_scratch_ptr[NA_MAX_DEVICES];
std::mutex na_mutex;
}  // namespace

extern "C" int cnf2_negshift_minima_noalloc(const double* negshift_dev, int32_t n_rows, int32_t Mc, double* minval_dev, int32_t* minstart_dev,
                                            void* stream)
{
	std::lock_guard<std::mutex> lock(na_mutex);
	if (n_rows < 0 || Mc <= 0 || (n_rows > 0 && (!negshift_dev || !minval_dev || !minstart_dev)))
	{
		g_create_error = "cnf2_negshift_minima_noalloc: null buffer or negative size";
		return CNF2_ERR_BAD_ARG;
	}
	if (n_rows == 0) return CNF2_OK;
	const int nchunk = (Mc + NA_CHUNK - 1) / NA_CHUNK;
	const size_t cells = (size_t)n_rows * nchunk;
	if (cells > NA_MAX_CELLS)
	{
		g_create_error = "cnf2_negshift_minima_noalloc: n_rows * ceil(Mc / 256) = " + std::to_string(cells) + " exceeds the static scratch of " +
		                 std::to_string(NA_MAX_CELLS) + " cells (8 MiB); split the rows over several calls or use cnf2_negshift_minima";
		return CNF2_ERR_BAD_ARG;
	}
	int dev = -1;
	cudaError_t err = cudaGetDevice(&dev);
	if (err != cudaSuccess || dev < 0 || dev >= NA_MAX_DEVICES)
	{
		g_create_error = std::string("cnf2_negshift_minima_noalloc: no current CUDA device (") + cudaGetErrorString(err) + ")";
		return CNF2_ERR_NO_DEVICE;
	}
	if (!na_scratch_ptr[dev])
	{
		double* p = nullptr;
		err = cudaGetSymbolAddress((void**)&p, na_chunkmin);
		if (err != cudaSuccess || !p)
		{
			g_create_error = std::string("cnf2_negshift_minima_noalloc: cudaGetSymbolAddress: ") + cudaGetErrorString(err);
			return CNF2_ERR_CUDA;
		}
		na_scratch_ptr[dev] = p;
	}
	double* const scratch = na_scratch_ptr[dev];
	cudaStream_t st = (cudaStream_t)stream;
	na_chunk_minima_kernel<<<(unsigned)((cells * 32 + 127) / 128), 128, 0, st>>>(negshift_dev, n_rows, Mc, nchunk, scratch);
	na_row_minima_kernel<<<(unsigned)(((size_t)n_rows * 32 + 127) / 128), 128, 0, st>>>(negshift_dev, n_rows, Mc, nchunk, scratch, minval_dev, minstart_dev);
	err = cudaGetLastError();
	if (err != cudaSuccess)
	{
		g_create_error = std::string("cnf2_negshift_minima_noalloc: launch: ") + cudaGetErrorString(err);
		return CNF2_ERR_CUDA;
	}
	return CNF2_OK;
}
