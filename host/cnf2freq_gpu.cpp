// cnf2freq_gpu -- host program of the B200 engine: the reference's command line, input formats, EM loop and outputs
// (row f4 of SURVEY.md section 8) around the C ABI of include/cnf2_engine.h.  Everything numerical happens behind that
// ABI on the GPU; this file is the part of cnF2freq's main (cnF2freq.cpp:7914-8198), of its readers
// (readalphamap :6669-6685, readalphaped :6495-6540, readalphadata :6542-6667, readgigidata :7450-7515) and of the print
// statements of doit (:5189-6410) that a drop-in has to keep:
//
//   cnf2freq_gpu --mapfile M --pedfile P --genfile G [--gigimapfile M --gigipedfile P] --count N [--limit L]
//                [--output FILE] [--tmppath DIR] [--deserialize FILE] [--parse-only] [--selfing]
//
// --selfing is the reference's compile-time knob SELFING (settings.h:14; lines listed with a generation above 2 are taken as
// selfed from generation 2 down, demo.sh:18-20) as a run-time switch: the reference is rebuilt for it, this program is not.
//
// --output FILE receives what the reference writes there: the state dump after every iteration (:8157-8192) and the last
// iteration's per-chromosome report; stdout carries the progress lines (likelihoods :5399-5401, "Inv:" :3450, "Scale
// factor now" :6392, "Correction at" / "Fixing point" of postmarkerdata).
//
// Several GPUs: one process per GPU, started N times with RANK / WORLD_SIZE / LOCAL_RANK in the environment (mpirun,
// torchrun --no-python, or a shell loop); the analysed individuals are dealt to the ranks in sibships and the statistics
// rows shared between ranks are summed by ncclAllReduce through the engine's hook (the reference's stale MPI reduce,
// :6241-6256).  The NCCL id travels through the file named by CNF2_NCCL_ID_FILE (default <tmppath>/cnf2_nccl_id).
#include <algorithm>
#include <cmath>
#include <cstdarg>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <map>
#include <string>
#include <vector>

#include "../include/cnf2_engine.h"

#ifdef CNF2_WITH_NCCL
#include <cuda_runtime.h>
#include <nccl.h>
#include <unistd.h>
#endif

namespace
{
constexpr int NUMPATHS = 128, NUMSHIFTS = 8, TURNBITS = 7;  // settings.h:32, 35, 40
constexpr int INDCOUNT = 1000000;                          // settings.h:9
constexpr uint8_t UNKNOWN = 0, SEXMARKER = 9;              // cnF2freq.cpp:225-226

struct Individ  // the fields of struct individ (:853-902) the host keeps per individual; per-marker data lives in flat arrays
{
	int n = 0;
	std::string name;
	int pars[2] = { 0, 0 };
	int gen = 0;
	bool empty = false, hasprior = false;
};

std::vector<double> markerposes;
std::vector<uint32_t> chromstarts;
std::vector<Individ> inds;  // index = individ::n; 0 = nobody
std::map<std::string, int> indmap;
std::vector<int> dous;
// per individual, per marker (row-major [n][marker]); grown as individuals appear
std::vector<uint8_t> markerdata, priordata;       // allele pairs
std::vector<double> markersure, priorsure, haploweight, relhaplo, negshift;

void die(const char* fmt, ...)
{
	va_list ap;
	va_start(ap, fmt);
	fprintf(stderr, "cnf2freq_gpu: ");
	vfprintf(stderr, fmt, ap);
	fprintf(stderr, "\n");
	va_end(ap);
	exit(1);
}

// getind(name) (:6478-6491): names are numbered in order of first appearance, "0" is nobody (zeroguy, :6493)
int getind(const std::string& name)
{
	auto it = indmap.find(name);
	if (it != indmap.end()) return it->second;
	const int n = (int)indmap.size();
	indmap[name] = n;
	if (n == 0) return 0;
	printf("Creating %d\n", n);
	const size_t M = markerposes.size();
	if ((int)inds.size() <= n) inds.resize(n + 1);
	inds[n].n = n;
	inds[n].name = name;
	markerdata.resize((size_t)(n + 1) * M * 2, UNKNOWN);
	priordata.resize((size_t)(n + 1) * M * 2, UNKNOWN);
	markersure.resize((size_t)(n + 1) * M * 2, 0.0);
	priorsure.resize((size_t)(n + 1) * M * 2, 0.0);
	haploweight.resize((size_t)(n + 1) * M, 0.5);  // :2494
	relhaplo.resize((size_t)(n + 1) * M, 0.5);     // :2496
	negshift.resize((size_t)(n + 1) * M, 0.0);
	return n;
}

void readalphamap(FILE* in)  // :6669-6685
{
	double prevval = 1e30, value;
	int nowmarker = 0;
	while (fscanf(in, "%lf", &value) == 1)
	{
		if (value < prevval) chromstarts.push_back(nowmarker);
		nowmarker++;
		markerposes.push_back(value);
		prevval = value;
	}
	chromstarts.push_back((uint32_t)markerposes.size());
}

void readalphaped(FILE* in)  // :6495-6540
{
	char me[255], father[255], mother[255];
	while (fscanf(in, "%254s %254s %254s", me, father, mother) == 3)
	{
		char line[255];
		if (!fgets(line, 255, in)) line[0] = 0;
		int gen = 0;
		if (sscanf(line, "%d", &gen) != 1) gen = 0;
		const int ime = getind(me), ff = getind(father), fm = getind(mother);
		// everybody is empty until genotypes arrive (:6511-6513)
		if (ime) inds[ime].empty = true;
		if (ff) inds[ff].empty = true;
		if (fm) inds[fm].empty = true;
		if (!ime) continue;
		if (gen >= 2 && ff && fm && !inds[ff].gen && !inds[fm].gen)
		{
			// an F2 listed with its grandparents: two implicit, empty F1 parents (:6515-6527)
			const int rp[2] = { getind(std::string(me) + "_aux_realf"), getind(std::string(me) + "_aux_realm") };
			for (int k = 0; k < 2; k++)
			{
				inds[ime].pars[k] = rp[k];
				inds[rp[k]].gen = 1;
				inds[rp[k]].pars[0] = ff;
				inds[rp[k]].pars[1] = fm;
				inds[rp[k]].empty = true;
			}
			inds[ime].gen = gen;
		}
		else
		{
			inds[ime].gen = gen;
			inds[ime].pars[0] = ff;
			inds[ime].pars[1] = fm;
		}
		if (gen >= 2) dous.push_back(ime);
	}
}

double binomial_pdf(int n, int k)  // boost::math::pdf(binomial_distribution<double>(n) [p = 0.5], k), :6599-6611
{
	return std::exp(std::lgamma(n + 1.0) - std::lgamma(k + 1.0) - std::lgamma(n - k + 1.0) + n * std::log(0.5));
}

void readalphadata(FILE* in)  // :6542-6667
{
	const size_t M = markerposes.size();
	const int haplo = getind("haplo");
	for (size_t x = 0; x < M; x++) markerdata[((size_t)haplo * M + x) * 2] = markerdata[((size_t)haplo * M + x) * 2 + 1] = SEXMARKER;
	char me[255];
	while (fscanf(in, "%254s", me) == 1)
	{
		const int ime = getind(me);
		if (!ime) die("genotypes for individual \"0\"");
		const bool doublehaplo = inds[ime].pars[1] == haplo;
		inds[ime].empty = false;
		for (size_t x = 0; x < M; x++)
		{
			char datastr[255];
			if (fscanf(in, "%254s", datastr) != 1) die("genotype file ends inside the row of %s", me);
			int data = 0, data2 = 0;
			const int numread = sscanf(datastr, "%d/%d", &data, &data2);
			uint8_t* md = &markerdata[((size_t)ime * M + x) * 2];
			double* ms = &markersure[((size_t)ime * M + x) * 2];
			haploweight[(size_t)ime * M + x] = 0.5;
			if (numread == 1)
			{
				switch (data)
				{
				case 0: md[0] = 1, md[1] = 1; break;
				case 1: md[0] = 1, md[1] = 2; break;
				case 2: md[0] = 2, md[1] = 2; break;
				default: md[0] = md[1] = UNKNOWN; break;
				}
				if (md[0] != UNKNOWN) ms[0] = ms[1] = 0.02;
			}
			else
			{
				printf("%s %d\n", datastr, numread);
				if (data == data2 && !data)
				{
					md[0] = md[1] = UNKNOWN;
					ms[0] = ms[1] = 0;
				}
				else
				{
					// read counts a/b: allele pair and error probabilities from a binomial mixture (:6599-6657)
					double sure1 = 0, sure2 = 0, probsum = 0;
					for (int rl1 = 0; rl1 <= data; rl1++)
						for (int rl2 = 0; rl2 <= data2; rl2++)
						{
							int l1 = rl1, l2 = rl2;
							double overallprob = (data ? binomial_pdf(data, l1) : 1) * (data2 ? binomial_pdf(data2, l2) : 1);
							double sureb1, sureb2;
							while (true)
							{
								sureb1 = 0.5;
								sureb2 = 0.5;
								if (l1 + l2) sureb1 = l1 / (double)(l1 + l2);
								if (data + data2 - l1 - l2) sureb2 = (data2 - l2) / (double)(data + data2 - l1 - l2);
								if (sureb1 + 1e-9 > 1 - sureb2) break;
								l1 = data - l1;
								l2 = data2 - l2;
							}
							overallprob *= std::pow(sureb1, l1) * std::pow(1 - sureb1, l2) * std::pow(sureb2, (data2 - l2)) * std::pow(1 - sureb2, data - l1);
							sure1 += sureb1 * overallprob;
							sure2 += sureb2 * overallprob;
							probsum += overallprob;
						}
					sure1 /= probsum;
					sure2 /= probsum;
					int mk[2] = { 2, 1 };
					double su[2] = { sure1, sure2 };
					for (int k = 0; k < 2; k++)  // an error rate over 0.5 is inverted
						if (su[k] > 0.5)
						{
							su[k] = 1 - su[k];
							mk[k] = k + 1;
						}
					md[0] = (uint8_t)mk[0], md[1] = (uint8_t)mk[1];
					ms[0] = su[0], ms[1] = su[1];
					printf("%d/%d turned into %d %d with %lf;%lf\n", data, data2, mk[0], mk[1], su[0], su[1]);
				}
				if (doublehaplo) md[1] = SEXMARKER;
			}
		}
		inds[ime].hasprior = true;  // priormarkerdata = markerdata (:6664-6665)
		std::copy(&markerdata[(size_t)ime * M * 2], &markerdata[(size_t)(ime + 1) * M * 2], &priordata[(size_t)ime * M * 2]);
		std::copy(&markersure[(size_t)ime * M * 2], &markersure[(size_t)(ime + 1) * M * 2], &priorsure[(size_t)ime * M * 2]);
	}
}

// readgigidata (:7450-7515) without Boost.Spirit: map lines `<int> <name> <cM> <int>` (one chromosome), ped lines
// `<family> <id> <father> <mother> <sex> <phenotype> (a b) x markers`; every individual is analysed (:7477)
void readgigidata(FILE* map, FILE* ped)
{
	chromstarts.push_back(0);
	char name[255];
	int a, b;
	double pos;
	while (fscanf(map, "%d %254s %lf %d", &a, name, &pos, &b) == 4) markerposes.push_back(pos);
	chromstarts.push_back((uint32_t)markerposes.size());
	const size_t M = markerposes.size();
	char fam[255], id[255], fa[255], mo[255], pheno[255];
	int sex;
	while (fscanf(ped, "%254s %254s %254s %254s %d %254s", fam, id, fa, mo, &sex, pheno) == 6)
	{
		const int ind = getind(id);
		if (!ind) die("gigi ped: individual \"0\"");
		dous.push_back(ind);
		inds[ind].pars[0] = getind(fa);
		inds[ind].pars[1] = getind(mo);
		for (size_t x = 0; x < M; x++)
		{
			int d1, d2;
			if (fscanf(ped, "%d %d", &d1, &d2) != 2) die("gigi ped: row of %s is short", id);
			markerdata[((size_t)ind * M + x) * 2] = (uint8_t)d1;
			markerdata[((size_t)ind * M + x) * 2 + 1] = (uint8_t)d2;
			markersure[((size_t)ind * M + x) * 2] = markersure[((size_t)ind * M + x) * 2 + 1] = 0;
		}
		inds[ind].hasprior = true;
		std::copy(&markerdata[(size_t)ind * M * 2], &markerdata[(size_t)(ind + 1) * M * 2], &priordata[(size_t)ind * M * 2]);
		std::copy(&markersure[(size_t)ind * M * 2], &markersure[(size_t)(ind + 1) * M * 2], &priorsure[(size_t)ind * M * 2]);
	}
}

// deserialize (:7757-7832): resume from the state dumps of an earlier run (its --output file): for every `<n> <name>` header
// of an individual known under that number and name, one line per marker `haploweight a b negshift sure1 sure2 ...`; later
// dumps of the file overwrite earlier ones.  Reports the phase switches per individual as the reference does.
void deserialize(FILE* in, const std::vector<int32_t>& children)
{
	const size_t M = markerposes.size();
	char line[1024];
	while (fgets(line, sizeof(line), in))
	{
		int n = 0, used = 0;
		char name[255];
		if (sscanf(line, "%d %254s %n", &n, name, &used) != 2 || line[used] != 0) continue;  // int_ >> word_ >> eoi
		auto it = indmap.find(name);
		if (n <= 0 || n >= (int)inds.size() || it == indmap.end() || it->second != n)
		{
			fprintf(stderr, "Supposed individual header not a header: %s", line);
			continue;
		}
		int oldphase = 0, switches = 0;
		for (size_t i = 0; i < M; i++)
		{
			double hw, ns, s1, s2;
			int a, b;
			if (!fgets(line, sizeof(line), in) || sscanf(line, "%lf %d %d %lf %lf %lf", &hw, &a, &b, &ns, &s1, &s2) != 6)
			{
				fprintf(stderr, "Reading haplotype for marker %zu for individual %s failed: %s", i, name, line);
				continue;
			}
			const size_t x = (size_t)n * M + i;
			haploweight[x] = hw;
			uint8_t* md = &markerdata[x * 2];
			bool inv = false, match = true;
			if (!(md[0] == a && md[1] == b))
			{
				if (!(md[0] == b && md[1] == a))
				{
					fprintf(stderr, "Genotype mismatch for marker %zu for individual %s (%d,%d) to  (%d,%d)\n", i, name, md[0], md[1], a, b);
					match = false;
				}
				else inv = true;
			}
			md[0] = (uint8_t)a, md[1] = (uint8_t)b;
			markersure[x * 2] = s1, markersure[x * 2 + 1] = s2;
			if (hw == 0.5 || a == b || !match) continue;
			const int newphase = 1 + ((hw > 0.5) ^ inv);
			if (oldphase && oldphase != newphase) switches++;
			oldphase = newphase;
		}
		const Individ& ind = inds[n];
		if (children[n] || (ind.pars[0] && !inds[ind.pars[0]].empty) || (ind.pars[1] && !inds[ind.pars[1]].empty))
			printf("Switches %d %s\t%d\n", n, name, switches);
	}
}

void log_line(void*, int32_t stream, const char* line)
{
	fputs(line, stream ? stderr : stdout);
}

#define CK(call)                                                                    \
	do                                                                              \
	{                                                                               \
		const int rc_ = (call);                                                     \
		if (rc_) die("%s failed (%d): %s", #call, rc_, cnf2_last_error(eng));       \
	} while (0)

#ifdef CNF2_WITH_NCCL
ncclComm_t g_comm, g_comm_control;  // one communicator per stream the engine calls the hook with (include/cnf2_engine.h)
void* g_control_stream = nullptr;
int nccl_hook(void*, double* buf, size_t count, void* stream)
{
	ncclComm_t comm = stream == g_control_stream ? g_comm_control : g_comm;
	return ncclAllReduce(buf, buf, count, ncclDouble, ncclSum, comm, (cudaStream_t)stream) == ncclSuccess ? 0 : 1;
}
#endif

// Sum of a host vector over the ranks (every rank contributes zeros for what it does not own): how rank 0 gets the rows
// of the other ranks' individuals for the likelihood lines and the state dump.
void sum_over_ranks(std::vector<double>& v, int world)
{
	if (world <= 1 || v.empty()) return;
#ifdef CNF2_WITH_NCCL
	static double* d = nullptr;
	static size_t cap = 0;
	if (cap < v.size())
	{
		if (d) cudaFree(d);
		if (cudaMalloc(&d, v.size() * sizeof(double)) != cudaSuccess) die("cudaMalloc for the gather failed");
		cap = v.size();
	}
	cudaMemcpy(d, v.data(), v.size() * sizeof(double), cudaMemcpyHostToDevice);
	if (ncclAllReduce(d, d, v.size(), ncclDouble, ncclSum, g_comm, 0) != ncclSuccess) die("ncclAllReduce failed");
	cudaStreamSynchronize(0);
	cudaMemcpy(v.data(), d, v.size() * sizeof(double), cudaMemcpyDeviceToHost);
#endif
}
}  // namespace

int main(int argc, char* argv[])
{
	std::string mapfn, pedfn, genfn, gigimap, gigiped, outfn, deserializefn, tmppath = ".";
	int COUNT = 3, limit = INDCOUNT;  // :7961-7962
	bool selfing = false;     // --selfing: settings.h:14 SELFING as a run-time switch
	bool parse_only = false;  // read the inputs, print what the readers made of them (the dump format of :8157-8192), stop: no GPU needed
	for (int i = 1; i < argc; i++)
	{
		const std::string a = argv[i];
		auto next = [&]() -> std::string {
			if (i + 1 >= argc) die("option %s needs a value", a.c_str());
			return argv[++i];
		};
		if (a == "--mapfile") mapfn = next();
		else if (a == "--pedfile") pedfn = next();
		else if (a == "--genfile") genfn = next();
		else if (a == "--gigimapfile") gigimap = next();
		else if (a == "--gigipedfile") gigiped = next();
		else if (a == "--output") outfn = next();
		else if (a == "--count") COUNT = atoi(next().c_str());
		else if (a == "--limit") limit = atoi(next().c_str());
		else if (a == "--tmppath") tmppath = next();
		else if (a == "--deserialize") deserializefn = next();
		else if (a == "--parse-only") parse_only = true;
		else if (a == "--selfing") selfing = true;
		else die("unknown option %s (supported: --mapfile --pedfile --genfile --gigimapfile --gigipedfile --count --limit --output --tmppath --deserialize --selfing)", a.c_str());
	}
	const int rank = getenv("RANK") ? atoi(getenv("RANK")) : 0, world = getenv("WORLD_SIZE") ? atoi(getenv("WORLD_SIZE")) : 1;
	const int local_rank = getenv("LOCAL_RANK") ? atoi(getenv("LOCAL_RANK")) : rank;
	if (rank && !freopen("/dev/null", "w", stdout)) die("cannot silence stdout");  // rank 0 speaks

	getind("0");  // zeroguy (:6493)
	if (!mapfn.empty())
	{
		FILE* f = fopen(mapfn.c_str(), "rt");
		if (!f) die("cannot open %s", mapfn.c_str());
		fprintf(stderr, "Reading map file %s\n", mapfn.c_str());
		readalphamap(f);
		fclose(f);
	}
	if (!pedfn.empty())
	{
		FILE* f = fopen(pedfn.c_str(), "rt");
		if (!f) die("cannot open %s", pedfn.c_str());
		fprintf(stderr, "Reading pedigree file %s\n", pedfn.c_str());
		readalphaped(f);
		fclose(f);
	}
	if (!genfn.empty())
	{
		FILE* f = fopen(genfn.c_str(), "rt");
		if (!f) die("cannot open %s", genfn.c_str());
		fprintf(stderr, "Reading genotype file %s\n", genfn.c_str());
		readalphadata(f);
		fclose(f);
	}
	if (!gigimap.empty() && !gigiped.empty())
	{
		FILE *fm = fopen(gigimap.c_str(), "rt"), *fp = fopen(gigiped.c_str(), "rt");
		if (!fm || !fp) die("cannot open the gigi files");
		readgigidata(fm, fp);
		fclose(fm), fclose(fp);
	}
	const size_t M = markerposes.size(), K = inds.size();
	const int C = (int)chromstarts.size() - 1;
	if (!M || K < 2 || C < 1) die("no markers or no individuals read");
	if ((size_t)limit < dous.size()) dous.resize(limit);  // :8122
	if (parse_only)
	{
		FILE* o = outfn.empty() ? stdout : fopen(outfn.c_str(), "w");
		if (!o) die("cannot open %s", outfn.c_str());
		fprintf(o, "markers %zu chromosomes %d individuals %zu analysed %zu\n", M, C, K - 1, dous.size());
		for (int c = 0; c <= C; c++) fprintf(o, "chromstart %u\n", chromstarts[c]);
		for (size_t i = 1; i < K; i++)
		{
			const bool analysed = std::find(dous.begin(), dous.end(), (int)i) != dous.end();
			fprintf(o, "%zu %s pars %d %d gen %d empty %d prior %d analysed %d\n", i, inds[i].name.c_str(), inds[i].pars[0], inds[i].pars[1], inds[i].gen,
			        (int)inds[i].empty, (int)inds[i].hasprior, (int)analysed);
			for (size_t j = 0; j < M; j++)
			{
				const size_t x = i * M + j;
				fprintf(o, "%f\t%d\t%d\t\t%f\t%lf %lf %lf\n", haploweight[x], markerdata[x * 2], markerdata[x * 2 + 1], negshift[x], markersure[x * 2],
				        markersure[x * 2 + 1], relhaplo[x]);
			}
		}
		if (o != stdout) fclose(o);
		return 0;
	}

	// ---- the engine ----
	cnf2_settings st;
	cnf2_default_settings(&st);
	st.device = local_rank;
	st.selfing = selfing ? 1 : 0;  // settings.h:14
	st.in_flight = C >= 3 ? 3 : 1;
	cnf2_engine* eng = nullptr;
	if (const int rc = cnf2_create(&eng, &st)) die("cnf2_create failed (%d): %s", rc, cnf2_last_error(nullptr));
	const double genrec[3] = { -1.0 / 50, -1.0 / 50, -1.0 / 50 };  // :7927-7943
	CK(cnf2_set_map(eng, markerposes.data(), (int32_t)M, chromstarts.data(), C, genrec));

	std::vector<int32_t> parents(K * 2, 0), gen(K, 0), descendants(K, 0), children(K, 0);
	std::vector<uint8_t> exists(K, 0), empty(K, 0), founder(K, 0), hasprior(K, 0);
	for (size_t i = 1; i < K; i++)
	{
		exists[i] = 1;
		parents[i * 2] = inds[i].pars[0], parents[i * 2 + 1] = inds[i].pars[1];
		gen[i] = inds[i].gen;
		empty[i] = inds[i].empty;
		hasprior[i] = inds[i].hasprior;
	}
	empty[0] = 1;
	// postmarkerdata (:8083-8085)
	CK(cnf2_preprocess(eng, (int32_t)K, parents.data(), exists.data(), empty.data(), markerdata.data(), markersure.data(), haploweight.data(),
	                   descendants.data(), children.data(), founder.data(), log_line, nullptr));

	if (!deserializefn.empty())  // :8087-8093
	{
		FILE* f = fopen(deserializefn.c_str(), "rt");
		if (!f) die("cannot open %s", deserializefn.c_str());
		printf("deserialize started.\n");
		deserialize(f, children);
		fclose(f);
		printf("deserialize finished.\n");
	}

	// analysed individuals of this rank: sibships dealt to the least loaded rank, the largest first (one rank: all of them)
	std::vector<int32_t> mine, shared;
	if (world > 1)
	{
		std::map<std::pair<int, int>, std::vector<int>> groups;
		for (int d : dous) groups[{ inds[d].pars[0], inds[d].pars[1] }].push_back(d);
		std::vector<std::pair<std::pair<int, int>, std::vector<int>>> order(groups.begin(), groups.end());
		std::stable_sort(order.begin(), order.end(), [](const auto& x, const auto& y) { return x.second.size() > y.second.size(); });
		std::vector<size_t> load(world, 0);
		std::vector<std::vector<int>> picks(world);
		const size_t fair = (dous.size() + world - 1) / world;
		for (auto& g : order)
			for (size_t i = 0; i < g.second.size(); i += fair)
			{
				const int r = (int)(std::min_element(load.begin(), load.end()) - load.begin());
				const size_t n = std::min(fair, g.second.size() - i);
				picks[r].insert(picks[r].end(), g.second.begin() + i, g.second.begin() + i + n);
				load[r] += n;
			}
		// rows more than one rank adds to: the all-reduce payload, the same list on every rank
		std::vector<int> seen(K, -1);
		std::vector<uint8_t> multi(K, 0);
		for (int r = 0; r < world; r++)
			for (int d : picks[r])
			{
				int tree[7] = { d, inds[d].pars[0], 0, 0, inds[d].pars[1], 0, 0 };
				for (int p = 0; p < 2; p++)
					if (const int q = inds[d].pars[p]) tree[1 + 3 * p + 1] = inds[q].pars[0], tree[1 + 3 * p + 2] = inds[q].pars[1];
				for (int t : tree)
					if (t)
					{
						if (seen[t] >= 0 && seen[t] != r) multi[t] = 1;
						seen[t] = r;
					}
			}
		for (size_t i = 1; i < K; i++)
			if (multi[i]) shared.push_back((int32_t)i);
		std::sort(picks[rank].begin(), picks[rank].end(), [&](int x, int y) {
			return std::find(dous.begin(), dous.end(), x) < std::find(dous.begin(), dous.end(), y);
		});
		mine.assign(picks[rank].begin(), picks[rank].end());
#ifdef CNF2_WITH_NCCL
		cudaSetDevice(local_rank);
		const std::string idfile = getenv("CNF2_NCCL_ID_FILE") ? getenv("CNF2_NCCL_ID_FILE") : tmppath + "/cnf2_nccl_id";
		ncclUniqueId id[2];  // statistics communicator, control communicator
		if (rank == 0)
		{
			if (ncclGetUniqueId(&id[0]) != ncclSuccess || ncclGetUniqueId(&id[1]) != ncclSuccess) die("ncclGetUniqueId failed");
			FILE* f = fopen((idfile + ".tmp").c_str(), "wb");
			if (!f || fwrite(id, sizeof(id), 1, f) != 1) die("cannot write %s", idfile.c_str());
			fclose(f);
			rename((idfile + ".tmp").c_str(), idfile.c_str());
		}
		else
		{
			FILE* f = nullptr;
			for (int tries = 0; tries < 600 && !(f = fopen(idfile.c_str(), "rb")); tries++) usleep(100000);
			if (!f || fread(id, sizeof(id), 1, f) != 1) die("cannot read the NCCL id from %s", idfile.c_str());
			fclose(f);
		}
		if (ncclCommInitRank(&g_comm, world, id[0], rank) != ncclSuccess) die("ncclCommInitRank failed");
		if (ncclCommInitRank(&g_comm_control, world, id[1], rank) != ncclSuccess) die("ncclCommInitRank (control) failed");
		if (rank == 0) remove(idfile.c_str());
		g_control_stream = cnf2_control_stream(eng);
		CK(cnf2_set_allreduce(eng, nccl_hook, nullptr));
#else
		die("built without NCCL (make -C host NCCL=1): WORLD_SIZE must be 1");
#endif
	}
	else mine.assign(dous.begin(), dous.end());
	std::vector<int32_t> alldous(dous.begin(), dous.end());
	CK(cnf2_set_shared_rows(eng, shared.data(), (int32_t)shared.size()));
	CK(cnf2_set_pedigree(eng, (int32_t)K, parents.data(), empty.data(), founder.data(), gen.data(), descendants.data(), mine.data(),
	                     (int32_t)mine.size(), alldous.data(), (int32_t)alldous.size()));
	for (int c = 0; c < C; c++) CK(cnf2_set_params(eng, c, markerdata.data(), markersure.data(), haploweight.data()));
	CK(cnf2_set_mstep_inputs(eng, relhaplo.data(), priordata.data(), priorsure.data(), hasprior.data(), children.data(), empty.data(),
	                         descendants.data()));
	CK(cnf2_set_em_state(eng, 0.013, 1.0, 0, 0));  // scalefactor, entropyfactor (:3573-3574)

	// fixtrees' masks of the analysed individuals for the likelihood lines (:5399)
	auto tree_masks = [&](int ind, int& flag2ignore, int& shiftignore) {
		flag2ignore = 1;
		shiftignore = 0;
		bool any1 = false;
		for (int lev1 = 0; lev1 < 2; lev1++)
		{
			const int l1 = inds[ind].pars[lev1];
			if (!l1) continue;
			const int flag2index = 1 + lev1 * 3, shiftval = 2 << lev1;
			if (!inds[l1].empty) flag2ignore |= 1 << flag2index;
			bool anypars = false;
			for (int lev2 = 0; lev2 < 2; lev2++)
			{
				const int l2 = inds[l1].pars[lev2];
				if (!l2 || inds[l2].empty) continue;
				flag2ignore |= 1 << (flag2index + lev2 + 1);
				anypars = true;
			}
			if (anypars) shiftignore |= shiftval;
			if (anypars || !inds[l1].empty) any1 = true;
		}
		if (any1) shiftignore |= 1;
		flag2ignore ^= NUMPATHS - 1;
		shiftignore ^= NUMSHIFTS - 1;
	};

	FILE* out = stdout;
	if (!outfn.empty() && rank == 0)
	{
		out = fopen(outfn.c_str(), "w");
		if (!out) die("cannot open %s", outfn.c_str());
	}
	std::vector<double> loglik(mine.size() * (size_t)C * 2, 0.0);
	// rows this rank reports in the gather: the ones only it touches, and (rank 0) the shared ones, which are equal everywhere
	std::vector<uint8_t> owned(K, 0);
	std::map<int, size_t> dou_pos;
	for (size_t j = 0; j < dous.size(); j++) dou_pos[dous[j]] = j;
	if (world > 1)
	{
		int32_t n_rows = 0;
		CK(cnf2_row_count(eng, &n_rows));
		std::vector<int32_t> rows(n_rows);
		CK(cnf2_row_ids(eng, rows.data()));
		for (int32_t id : rows) owned[id] = 1;
		if (rank)
			for (int32_t id : shared) owned[id] = 0;
	}
	std::vector<int32_t> inv_ids(4096), inv_markers(4096);
	for (int it = 0; it < COUNT; it++)
	{
		const bool early = it < 1;  // :8130
		FILE* rep = (it == COUNT - 1) ? out : stdout;
		if (!early)
		{
			// doit (:5189-6410).  individ::children is recounted over the analysed individuals by every call (:5226, :5246-5259)
			std::fill(children.begin(), children.end(), 0);
			for (int d : dous)
				for (int p = 0; p < 2; p++)
					if (const int q = inds[d].pars[p]) children[q]++;
			CK(cnf2_set_mstep_rows(eng, hasprior.data(), children.data(), empty.data(), descendants.data()));
			CK(cnf2_begin_iteration(eng));
			const int ahead = C >= 3 ? 2 : 0;  // E-steps of the next chromosomes run beside the block behind this one
			for (int c = 0; c < std::min(ahead, C); c++) CK(cnf2_estep(eng, c));
			for (int c = 0; c < C; c++)
			{
				if (c + ahead < C || ahead == 0) CK(cnf2_estep(eng, c + ahead));
				CK(cnf2_get_stats(eng, c, nullptr, nullptr, nullptr, nullptr, loglik.data()));
				std::vector<double> ll(dous.size() * 2, 0.0);  // in the order of the whole run's analysed individuals
				for (size_t j = 0; j < mine.size(); j++)
				{
					const size_t at = world > 1 ? dou_pos[mine[j]] : j;
					ll[at * 2] = loglik[(j * C + c) * 2], ll[at * 2 + 1] = loglik[(j * C + c) * 2 + 1];
				}
				sum_over_ranks(ll, world);
				for (size_t j = 0; j < dous.size(); j++)
				{
					int f2i, shi;
					tree_masks(dous[j], f2i, shi);
					const int shiftend = inds[dous[j]].gen < 2 ? 2 : NUMSHIFTS;  // :5359
					printf("%d,%03d,%03d: %lf\t%lf %d\n", dous[j], f2i, shi, ll[j * 2], ll[j * 2 + 1], shiftend);
				}
				for (int d : dous) fprintf(rep, "%s:%d\n\n", inds[d].name.c_str(), c + 1);  // :6183-6188 (the rows are empty at HEAD, :5873)
				fflush(rep);
				for (size_t i = 1; i < K; i++) fprintf(rep, "FIRST PASS: %d\n", (int)i);  // :6259
				int32_t n_inv = 0;
				CK(cnf2_em_pass(eng, c, inv_ids.data(), inv_markers.data(), (int32_t)inv_ids.size(), &n_inv));
				for (int k = 0; k < std::min<int32_t>(n_inv, (int32_t)inv_ids.size()); k++) printf("Inv: %d %d\n", inv_ids[k], inv_markers[k]);
				for (size_t i = 1; i < K; i++) fprintf(rep, "SKEWNESS PASS: %d\n", (int)i);  // :6351
				double sf, ent;
				int32_t o1, o2;
				CK(cnf2_get_em_state(eng, &sf, &ent, &o1, &o2, nullptr));
				printf("Scale factor now %lf, entropy %lf, hitnnn %d\n", sf, ent, o1);  // :6392
				CK(cnf2_get_negshift(eng, c, negshift.data()));  // the chromosome's statistics block is reused by a later E-step
			}
			for (int c = 0; c < C; c++) CK(cnf2_get_params(eng, c, markerdata.data(), markersure.data(), haploweight.data()));
			if (world > 1)
			{
				// every rank ends the iteration with the whole state: six values per (individual, marker), zeros where not owned
				std::vector<double> pack(K * M * 6, 0.0);
				for (size_t i = 1; i < K; i++)
					if (owned[i])
						for (size_t j = 0; j < M; j++)
						{
							const size_t x = i * M + j;
							double* q = &pack[x * 6];
							q[0] = haploweight[x], q[1] = markerdata[x * 2], q[2] = markerdata[x * 2 + 1], q[3] = markersure[x * 2], q[4] = markersure[x * 2 + 1];
							q[5] = negshift[x];
						}
				std::vector<double> touched(K, 0.0);
				for (size_t i = 1; i < K; i++) touched[i] = owned[i];
				sum_over_ranks(pack, world);
				sum_over_ranks(touched, world);
				for (size_t i = 1; i < K; i++)
					if (touched[i] > 0)
						for (size_t j = 0; j < M; j++)
						{
							const size_t x = i * M + j;
							const double* q = &pack[x * 6];
							haploweight[x] = q[0], markerdata[x * 2] = (uint8_t)q[1], markerdata[x * 2 + 1] = (uint8_t)q[2], markersure[x * 2] = q[3];
							markersure[x * 2 + 1] = q[4], negshift[x] = q[5];
						}
			}
		}
		fflush(stdout);
		fflush(out);
		// the state after every iteration (:8157-8192)
		if (rank == 0)
			for (size_t i2 = 1; i2 < K; i2++)
			{
				if ((long)i2 > (long)limit) continue;
				fprintf(out, "%d %s\n", (int)i2, inds[i2].name.c_str());
				for (size_t j = 0; j < M; j++)
				{
					const size_t x = i2 * M + j;
					if (inds[i2].hasprior)
						fprintf(out, "%f\t%d\t%d\t\t%f\t%lf %lf %lf\t%d\t%d\t%lf\t%lf\n", haploweight[x], markerdata[x * 2], markerdata[x * 2 + 1], negshift[x],
						        markersure[x * 2], markersure[x * 2 + 1], relhaplo[x], priordata[x * 2], priordata[x * 2 + 1], priorsure[x * 2], priorsure[x * 2 + 1]);
					else
						fprintf(out, "%f\t%d\t%d\t\t%f\t%lf %lf %lf\n", haploweight[x], markerdata[x * 2], markerdata[x * 2 + 1], negshift[x], markersure[x * 2],
						        markersure[x * 2 + 1], relhaplo[x]);
					negshift[x] = 0;
				}
			}
		fflush(stdout);
		fflush(out);
	}
	if (out != stdout) fclose(out);
	cnf2_destroy(eng);
	return 0;
}
