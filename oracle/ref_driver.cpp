// TEST INFRASTRUCTURE ONLY -- not part of the product.
//
// Driver appended (by oracle/Makefile) to lines 1-6794 of the UNMODIFIED reference translation unit
// /root/reference/cnF2freq.cpp, forming the binary oracle/_ref/cnf2ref.  The reference text is piped to
// the compiler straight from /root/reference; no reference source is stored in this repository.
//
// What this file adds (all of it is ours):
//   * main(): the reference main()'s flow (cnF2freq.cpp:7927-7943 genrec setup, readalpha* calls :7990-8009,
//     postmarkerdata :8083-8085, the EM loop :8128-8136) without boost::program_options;
//   * an fscanf restatement of readgigidata (cnF2freq.cpp:7450-7515; the original needs Boost.Spirit);
//   * observation hooks that do NOT modify the reference code path: doit() is handed a fopencookie()
//     FILE* as `out`, and stderr is swapped for another cookie stream.  The reference itself prints
//     "LAST: <n> <name>" to stderr after finishing each analysed individual (cnF2freq.cpp:5908) and
//     "FIRST PASS: <i>" to `out` after the E-step of a chromosome and before its M-step
//     (cnF2freq.cpp:6259).  On those lines we dump the thread-local forward-backward arrays
//     (fwbw/fwbwfactors, cnF2freq.cpp:392-394) and the EM sufficient statistics
//     (haplobase/haplocount/infprobs/homozyg, cnF2freq.cpp:895-902) exactly as the reference computed them.
//   * binary dumps of the problem and of the parameters before/after each EM iteration;
//   * a --mode time leg used as the CPU baseline (bench.py --impl reference).
//
// Dump format ("CNF2DUMP"): repeated records  "ARRAY <name> <dtype> <ndim> <d0> ... <dn-1>\n" + raw
// little-endian bytes + "\n".  dtype in {f64,i32,u8}.  Reader: tests/cnf2dump.py.

#include <unistd.h>
#include <fcntl.h>

namespace cnf2drv
{
static FILE* realstderr = nullptr;
static std::string dumpprefix;
static std::string statedumpfn;  // the per-iteration text dump of cnF2freq.cpp:8157-8192 (what main writes to --output)
static bool dumpfwbw = false;
static bool quiet = false;
static int curiter = 0;
static int curchrom = 0;
static bool statsdumpedforchrom = false;
static int maxn = 0;

static void writearray(FILE* f, const char* name, const char* dtype, std::vector<size_t> dims, const void* data, size_t elsize)
{
	size_t total = 1;
	fprintf(f, "ARRAY %s %s %zu", name, dtype, dims.size());
	for (size_t d : dims)
	{
		fprintf(f, " %zu", d);
		total *= d;
	}
	fprintf(f, "\n");
	if (total) fwrite(data, elsize, total, f);
	fprintf(f, "\n");
}

static void wf64(FILE* f, const char* name, std::vector<size_t> dims, const std::vector<double>& v)
{
	writearray(f, name, "f64", dims, v.data(), 8);
}
static void wi32(FILE* f, const char* name, std::vector<size_t> dims, const std::vector<int>& v)
{
	writearray(f, name, "i32", dims, v.data(), 4);
}
static void wu8(FILE* f, const char* name, std::vector<size_t> dims, const std::vector<unsigned char>& v)
{
	writearray(f, name, "u8", dims, v.data(), 1);
}

static int computemaxn()
{
	int m = 0;
	for (int i = 1; i < INDCOUNT; i++)
	{
		if (individer[i]) m = i;
	}
	return m;
}

static void dumpproblem()
{
	if (dumpprefix.empty()) return;
	FILE* f = fopen((dumpprefix + ".problem").c_str(), "wb");
	size_t K = maxn + 1;
	size_t M = markerposes.size();
	wf64(f, "markerposes", { M }, markerposes);
	std::vector<int> cs(chromstarts.begin(), chromstarts.end());
	wi32(f, "chromstarts", { cs.size() }, cs);
	std::vector<double> gr(genrec, genrec + 3);
	wf64(f, "genrec", { 3 }, gr);
	std::vector<unsigned char> exists(K), empty(K), founder(K), hasprior(K);
	std::vector<int> gen(K), par0(K), par1(K), desc(K), children(K);
	std::string names;
	for (size_t i = 0; i < K; i++)
	{
		individ* ind = i ? individer[i] : nullptr;
		exists[i] = ind != nullptr;
		if (ind)
		{
			empty[i] = ind->empty;
			founder[i] = ind->founder;
			gen[i] = ind->gen;
			par0[i] = ind->pars[0] ? ind->pars[0]->n : 0;
			par1[i] = ind->pars[1] ? ind->pars[1]->n : 0;
			desc[i] = ind->descendants;
			children[i] = ind->children;
			hasprior[i] = ind->priormarkerdata.size() > 0;
			names += ind->name;
		}
		names += "\n";
	}
	wu8(f, "exists", { K }, exists);
	wu8(f, "empty", { K }, empty);
	wu8(f, "founder", { K }, founder);
	wu8(f, "hasprior", { K }, hasprior);
	wi32(f, "gen", { K }, gen);
	wi32(f, "par0", { K }, par0);
	wi32(f, "par1", { K }, par1);
	wi32(f, "descendants", { K }, desc);
	wi32(f, "children", { K }, children);
	std::vector<unsigned char> nm(names.begin(), names.end());
	wu8(f, "names", { nm.size() }, nm);
	std::vector<int> d;
	for (individ* ind : dous) d.push_back(ind->n);
	wi32(f, "dous", { d.size() }, d);
	fclose(f);
}

// Parameters that drive the E-step (and the M-step's starting point).
static void dumpstate(const std::string& suffix)
{
	if (dumpprefix.empty()) return;
	FILE* f = fopen((dumpprefix + suffix).c_str(), "wb");
	size_t K = maxn + 1;
	size_t M = markerposes.size();
	std::vector<unsigned char> alleles(K * M * 2), prioralleles(K * M * 2), founder(K);
	std::vector<double> sure(K * M * 2), priorsure(K * M * 2), hw(K * M), rh(K * M), ns(K * M);
	std::vector<int> children(K), lastinved;
	for (size_t i = 1; i < K; i++)
	{
		individ* ind = individer[i];
		if (!ind) continue;
		founder[i] = ind->founder;
		children[i] = ind->children;
		for (size_t m = 0; m < M && m < ind->markerdata.size(); m++)
		{
			alleles[(i * M + m) * 2] = ind->markerdata[m].first.value();
			alleles[(i * M + m) * 2 + 1] = ind->markerdata[m].second.value();
			sure[(i * M + m) * 2] = ind->markersure[m].first;
			sure[(i * M + m) * 2 + 1] = ind->markersure[m].second;
			hw[i * M + m] = ind->haploweight[m];
			rh[i * M + m] = ind->relhaplo[m];
			ns[i * M + m] = ind->negshift[m];
			if (ind->priormarkerdata.size() > m)
			{
				prioralleles[(i * M + m) * 2] = ind->priormarkerdata[m].first.value();
				prioralleles[(i * M + m) * 2 + 1] = ind->priormarkerdata[m].second.value();
				priorsure[(i * M + m) * 2] = ind->priormarkersure[m].first;
				priorsure[(i * M + m) * 2 + 1] = ind->priormarkersure[m].second;
			}
		}
	}
	wu8(f, "alleles", { K, M, 2 }, alleles);
	wf64(f, "sure", { K, M, 2 }, sure);
	wf64(f, "haploweight", { K, M }, hw);
	wf64(f, "relhaplo", { K, M }, rh);
	wf64(f, "negshift", { K, M }, ns);
	wu8(f, "prioralleles", { K, M, 2 }, prioralleles);
	wf64(f, "priorsure", { K, M, 2 }, priorsure);
	wu8(f, "founder", { K }, founder);
	wi32(f, "children", { K }, children);
	std::vector<double> sc = { scalefactor, entropyfactor, (double)oldhitnnn, (double)oldhitnnn2 };
	wf64(f, "scalars", { 4 }, sc);
	fclose(f);
}

// EM sufficient statistics as left by the E-step of one chromosome (cnF2freq.cpp:5875-5904), read
// before the M-step touches them.
static void dumpstats()
{
	if (dumpprefix.empty()) return;
	char suffix[64];
	sprintf(suffix, ".iter%d.chrom%d.stats", curiter, curchrom);
	FILE* f = fopen((dumpprefix + suffix).c_str(), "wb");
	size_t K = maxn + 1;
	size_t M = markerposes.size();
	std::vector<double> hb(K * M), hc(K * M), ip(K * M * 4), hz(K * M * 2);
	std::vector<int> oddkeys(1);
	for (size_t i = 1; i < K; i++)
	{
		individ* ind = individer[i];
		if (!ind || !ind->haplocount.size()) continue;
		for (size_t m = 0; m < M; m++)
		{
			hb[i * M + m] = ind->haplobase[m];
			hc[i * M + m] = ind->haplocount[m];
			for (int side = 0; side < 2; side++)
			{
				for (auto& kv : ind->infprobs[m][side])
				{
					int a = kv.first.value();
					if (a == 1 || a == 2) ip[((i * M + m) * 2 + side) * 2 + (a - 1)] = kv.second;
					else oddkeys[0]++;
				}
			}
			hz[(i * M + m) * 2] = ind->homozyg[m][0];
			hz[(i * M + m) * 2 + 1] = ind->homozyg[m][1];
		}
	}
	wf64(f, "haplobase", { K, M }, hb);
	wf64(f, "haplocount", { K, M }, hc);
	wf64(f, "infprobs", { K, M, 2, 2 }, ip);
	wf64(f, "homozyg", { K, M, 2 }, hz);
	wi32(f, "oddkeys", { 1 }, oddkeys);
	fclose(f);
}

static FILE* fwbwfile = nullptr;

// Called when the reference reports "LAST: n name" (cnF2freq.cpp:5908): the thread-local fwbw arrays
// still hold this individual's sweeps for the current chromosome.  Single-threaded runs only.
static void dumpfwbwfor(int n)
{
	if (dumpprefix.empty() || !dumpfwbw) return;
	if (!fwbwfile)
	{
		char suffix[64];
		sprintf(suffix, ".iter%d.fwbw", curiter);
		fwbwfile = fopen((dumpprefix + suffix).c_str(), "wb");
	}
	size_t c0 = chromstarts[curchrom], c1 = chromstarts[curchrom + 1];
	size_t Mc = c1 - c0;
	std::vector<double> a(NUMSHIFTS * Mc * 3 * NUMTYPES), fac(NUMSHIFTS * Mc * 3);
	std::vector<int> mask(NUMSHIFTS);
	for (unsigned int s = 0; s < NUMSHIFTS; s++)
	{
		// generation was already incremented for this individual (cnF2freq.cpp:5302)
		mask[s] = (fwbwdone[s] >> 2) == generation ? (fwbwdone[s] & 3) : 0;
		if (!mask[s]) continue;
		for (size_t m = 0; m < Mc; m++)
		{
			for (int pad = 0; pad < 3; pad++)
			{
				for (int k = 0; k < NUMTYPES; k++)
				{
					a[((s * Mc + m) * 3 + pad) * NUMTYPES + k] = fwbw[s][c0 + m][pad][k];
				}
				fac[(s * Mc + m) * 3 + pad] = fwbwfactors[s][c0 + m][pad];
			}
		}
	}
	char name[64];
	sprintf(name, "fwbw_c%d_n%d", curchrom, n);
	wf64(fwbwfile, name, { NUMSHIFTS, Mc, 3, (size_t)NUMTYPES }, a);
	sprintf(name, "fwbwfactors_c%d_n%d", curchrom, n);
	wf64(fwbwfile, name, { NUMSHIFTS, Mc, 3 }, fac);
	sprintf(name, "fwbwmask_c%d_n%d", curchrom, n);
	wi32(fwbwfile, name, { NUMSHIFTS }, mask);
}

struct linebuf
{
	std::string partial;
	void (*online)(const std::string&);
	FILE* forward;
};

static ssize_t cookiewrite(void* c, const char* buf, size_t size)
{
	linebuf* lb = (linebuf*)c;
	if (lb->forward && !quiet) fwrite(buf, 1, size, lb->forward);
	for (size_t i = 0; i < size; i++)
	{
		if (buf[i] == '\n')
		{
			lb->online(lb->partial);
			lb->partial.clear();
		}
		else lb->partial.push_back(buf[i]);
	}
	return size;
}

static void onerrline(const std::string& line)
{
	int n;
	if (sscanf(line.c_str(), "LAST: %d", &n) == 1)
	{
		// a new E-step individual; if the previous chromosome's stats were dumped we have moved on
		if (statsdumpedforchrom)
		{
			curchrom++;
			statsdumpedforchrom = false;
		}
		dumpfwbwfor(n);
	}
}

static void onoutline(const std::string& line)
{
	int n;
	if (sscanf(line.c_str(), "FIRST PASS: %d", &n) == 1)
	{
		if (!statsdumpedforchrom)
		{
			dumpstats();
			statsdumpedforchrom = true;
		}
	}
}

static FILE* makecookie(void (*online)(const std::string&), FILE* forward)
{
	linebuf* lb = new linebuf{ "", online, forward };
	cookie_io_functions_t io = { nullptr, cookiewrite, nullptr, nullptr };
	FILE* f = fopencookie(lb, "w", io);
	setvbuf(f, nullptr, _IONBF, 0);
	return f;
}

// fscanf restatement of readgigidata (cnF2freq.cpp:7450-7515).  Map line: <int> <name> <cM> <int>;
// ped line: <fam> <id> <father> <mother> <sex> <pheno> then one allele pair per marker.
static void readgigi(const char* mapfn, const char* pedfn)
{
	FILE* map = fopen(mapfn, "rt");
	FILE* ped = fopen(pedfn, "rt");
	if (!map || !ped)
	{
		fprintf(stderr, "cannot open gigi files\n");
		exit(2);
	}
	chromstarts.push_back(0);
	int dummy1, dummy2;
	char name[256];
	double pos;
	while (fscanf(map, "%d %255s %lf %d", &dummy1, name, &pos, &dummy2) == 4)
	{
		markernames[name] = markerposes.size();
		markerposes.push_back(pos);
	}
	chromstarts.push_back(markerposes.size());

	char fam[256], id[256], fa[256], mo[256], pheno[256];
	int sex;
	while (fscanf(ped, "%255s %255s %255s %255s %d %255s", fam, id, fa, mo, &sex, pheno) == 6)
	{
		individ* ind = getind(std::string(id));
		dous.push_back(ind);
		ind->pars[0] = getind(std::string(fa));
		ind->pars[1] = getind(std::string(mo));
		ind->sex = sex - 1;
		for (size_t m = 0; m < markerposes.size(); m++)
		{
			int a, b;
			if (fscanf(ped, "%d %d", &a, &b) != 2)
			{
				fprintf(stderr, "short gigi ped line\n");
				exit(2);
			}
			ind->markerdata[m] = make_pair(a * MarkerValue, b * MarkerValue);
			ind->markersure[m] = make_pair(0, 0);
		}
		ind->priormarkerdata = ind->markerdata;
		ind->priormarkersure = ind->markersure;
	}
	fclose(map);
	fclose(ped);
}
} // namespace cnf2drv

int main(int argc, char* argv[])
{
	using namespace cnf2drv;
	// cnF2freq.cpp:7927-7943
	discstep = 1;
	sexc = 2;
	baserec[0] = -discstep / 50.0;
	baserec[1] = baserec[0];
	for (int gen = 0; gen < 2; gen++) genrec[gen] = baserec[0];
	genrec[2] = baserec[0];

	std::string mapfn, pedfn, genfn, gigimap, gigiped, outfn, mode = "doit";
	int COUNT = 3;
	int limit = INDCOUNT;
	for (int i = 1; i < argc; i++)
	{
		std::string a = argv[i];
		auto next = [&]() -> std::string { return i + 1 < argc ? argv[++i] : ""; };
		if (a == "--mapfile") mapfn = next();
		else if (a == "--pedfile") pedfn = next();
		else if (a == "--genfile") genfn = next();
		else if (a == "--gigimapfile") gigimap = next();
		else if (a == "--gigipedfile") gigiped = next();
		else if (a == "--output") outfn = next();
		else if (a == "--count") COUNT = atoi(next().c_str());
		else if (a == "--limit") limit = atoi(next().c_str());
		else if (a == "--dump") dumpprefix = next();
		else if (a == "--statedump") statedumpfn = next();
		else if (a == "--dumpfwbw") dumpfwbw = true;
		else if (a == "--quiet") quiet = true;
		else if (a == "--mode") mode = next();
		else
		{
			fprintf(stderr, "unknown option %s\n", a.c_str());
			return 2;
		}
	}

	int resultfd = -1;
	if (mode == "time")
	{
		// the reference's own progress printing goes to /dev/null; results go to the original stdout
		resultfd = dup(1);
		fflush(stdout);
		freopen("/dev/null", "w", stdout);
		freopen("/dev/null", "w", stderr);
	}

	if (mapfn != "")
	{
		FILE* f = fopen(mapfn.c_str(), "rt");
		if (!f) return 2;
		readalphamap(f);
		fclose(f);
	}
	if (pedfn != "")
	{
		FILE* f = fopen(pedfn.c_str(), "rt");
		if (!f) return 2;
		readalphaped(f);
		fclose(f);
	}
	if (genfn != "")
	{
		FILE* f = fopen(genfn.c_str(), "rt");
		if (!f) return 2;
		readalphadata(f);
		fclose(f);
	}
	if (gigimap != "" && gigiped != "") readgigi(gigimap.c_str(), gigiped.c_str());

	double t0 = omp_get_wtime();
	// cnF2freq.cpp:8083-8085
	CORRECTIONINFERENCE = true;
	postmarkerdata(limit);
	CORRECTIONINFERENCE = false;
	double tpost = omp_get_wtime() - t0;

	if (limit < (int)dous.size()) dous.resize(limit);
	maxn = computemaxn();
	dumpproblem();

	FILE* out = stdout;
	if (mode == "doit")
	{
		realstderr = stderr;
		out = makecookie(onoutline, outfn != "" ? fopen(outfn.c_str(), "w") : (quiet ? nullptr : stdout));
		stderr = makecookie(onerrline, realstderr);
	}
	else if (outfn != "")
	{
		out = fopen(outfn.c_str(), "w");
	}

	size_t analysed = dous.size();
	size_t M = markerposes.size();
	// cnF2freq.cpp:8128-8136: iteration 0 is "early" and only dumps
	for (int i = 0; i < COUNT; i++)
	{
		early = (i < 1);
		curiter = i;
		curchrom = 0;
		statsdumpedforchrom = false;
		if (!early)
		{
			char suffix[64];
			sprintf(suffix, ".iter%d.params", i);
			dumpstate(suffix);
			double t1 = omp_get_wtime();
			doit<false, genotypereporter>(out, true);
			double dt = omp_get_wtime() - t1;
			if (fwbwfile)
			{
				fclose(fwbwfile);
				fwbwfile = nullptr;
			}
			sprintf(suffix, ".iter%d.state", i);
			dumpstate(suffix);
			if (mode == "time")
			{
				char line[256];
				int len = sprintf(line, "TIME iter %d seconds %.6f individuals %zu markers %zu states %d threads %d\n", i, dt, analysed, M, NUMTYPES, omp_get_max_threads());
				if (write(resultfd, line, len) < 0) return 3;
			}
		}
		// cnF2freq.cpp:8157-8192: the state after every iteration, in main's formats
		if (!statedumpfn.empty())
		{
			static FILE* sd = fopen(statedumpfn.c_str(), "w");
			for (int i2 = 1; i2 <= maxn; i2++)
			{
				if (i2 > limit) continue;
				individ* ind = individer[i2];
				if (!ind || !ind->haplocount.size()) continue;
				fprintf(sd, "%d %s\n", i2, ind->name.c_str());
				for (unsigned int c = 0; c < chromstarts.size() - 1; c++)
					for (unsigned int j = chromstarts[c]; j < chromstarts[c + 1]; j++)
					{
						if (ind->priormarkerdata.size() > j)
							fprintf(sd, "%f\t%d\t%d\t\t%f\t%lf %lf %lf\t%d\t%d\t%lf\t%lf\n", ind->haploweight[j], ind->markerdata[j].first.value(),
							        ind->markerdata[j].second.value(), ind->negshift[j], ind->markersure[j].first, ind->markersure[j].second,
							        RELSKEWS ? ind->relhaplo[j] : 0, ind->priormarkerdata[j].first.value(), ind->priormarkerdata[j].second.value(),
							        ind->priormarkersure[j].first, ind->priormarkersure[j].second);
						else
							fprintf(sd, "%f\t%d\t%d\t\t%f\t%lf %lf %lf\n", ind->haploweight[j], ind->markerdata[j].first.value(),
							        ind->markerdata[j].second.value(), ind->negshift[j], ind->markersure[j].first, ind->markersure[j].second,
							        RELSKEWS ? ind->relhaplo[j] : 0);
					}
			}
			fflush(sd);
		}
		// cnF2freq.cpp:8187: negshift is reset after each iteration's dump
		for (int i2 = 1; i2 <= maxn; i2++)
		{
			individ* ind = individer[i2];
			if (!ind) continue;
			for (size_t j = 0; j < ind->negshift.size(); j++) ind->negshift[j] = 0;
		}
	}
	if (mode == "time")
	{
		char line[256];
		int len = sprintf(line, "TIME postmarkerdata seconds %.6f\n", tpost);
		if (write(resultfd, line, len) < 0) return 3;
	}
	fflush(stdout);
	return 0;
}
