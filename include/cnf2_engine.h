/*
 * cnf2_engine.h -- C ABI of the B200 (sm_100a) engine for the hot path of cnettel/cnF2freq.
 *
 * The reference has no plugin/FFI seam (one translation unit); the boundary is cut
 * inside doit(): the body of the OpenMP loop over analysed individuals for one chromosome,
 * cnF2freq.cpp:5294-5907.  What the reference computes there -- the 8 shift-mode forward/backward sweeps
 * (initfwbw :2074-2120, realanalyze :2145-2418, adjustprobs :1579-1670, trackpossible :1075-1359), the
 * marker x state x shift x orientation posterior enumeration (:5406-5583, ignoreflag2 :3462-3496) and the
 * transfer into the per-individual EM sufficient statistics (moveinfprobs/movehaplos :3577-3616,
 * :5875-5904) -- is what cnf2_estep() computes on the GPU.  Everything is fp64, like the reference.
 *
 * Conventions
 *   - individual ids are 1..n_ind-1, id 0 means "no individual" (individ::n / individer[], :2447-2450);
 *   - all [n_ind][n_markers] arrays are row-major over ALL markers of ALL chromosomes (the reference's
 *     per-individual vectors, :853-902), caller-owned; the engine copies what it needs on set_* and
 *     fills caller buffers on get_*;
 *   - every call returns 0 on success or a negative cnf2_status; cnf2_last_error() has the text;
 *   - one host thread per engine; several engines (one per GPU / rank) may coexist;
 *   - there is NO CPU fallback: without a CUDA device cnf2_create() fails with CNF2_ERR_NO_DEVICE.
 *
 * Reference-side binding: see INTEGRATION.md (the replacement for the loop at cnF2freq.cpp:5294).
 */
#ifndef CNF2_ENGINE_H
#define CNF2_ENGINE_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct cnf2_engine cnf2_engine; /* opaque; owns all device memory */

typedef enum cnf2_status
{
	CNF2_OK = 0,
	CNF2_ERR_NO_DEVICE = -1,  /* no CUDA device / driver: the engine never falls back to the CPU */
	CNF2_ERR_BAD_ARG = -2,
	CNF2_ERR_CUDA = -3,       /* a CUDA runtime call or kernel failed; see cnf2_last_error() */
	CNF2_ERR_STATE = -4,      /* calls out of order (e.g. estep before set_params) */
	CNF2_ERR_UNSUPPORTED = -5 /* a settings.h combination the kernels are not built for */
} cnf2_status;

/* settings.h knobs that shape the state space (settings.h:3-47).  The kernels are instantiated for the
 * HEAD configuration: NUMGEN 3 (64 states, 128 orientations, 8 shift modes), SELFING false,
 * RELSKEWSTATES false, HAPLOTYPING true, DOINFPROBS true -- and for SELFING true (settings.h:7, 14, 21-27, 46: 256
 * states = 64 x the value of two selfing bits, 192 valid; the PlantImpute mode of demo.sh:18-20, lines listed with a
 * generation above 2).  In the reference SELFING is a compile-time constant; here it is chosen per engine.  Under SELFING
 * every unit runs through the literal kernel (cnf2::estep_kernel<true>; the warp kernel is built for 64 states), every
 * analysed individual needs generation 2..32 (1 << (gen - 2) of cnF2freq.cpp:2095, 2321), and cnf2_preprocess scans the
 * 256 states.  Anything else (NUMGEN != 3, RELSKEWSTATES, !HAPLOTYPING, !DOINFPROBS) -> CNF2_ERR_UNSUPPORTED. */
typedef struct cnf2_settings
{
	int32_t numgen;        /* settings.h:19  NUMGEN */
	int32_t selfing;       /* settings.h:14  SELFING: 0 or 1 */
	int32_t relskewstates; /* settings.h:16  RELSKEWSTATES */
	int32_t haplotyping;   /* settings.h:36  HAPLOTYPING */
	int32_t doinfprobs;    /* settings.h:12  DOINFPROBS */
	int32_t device;        /* CUDA device ordinal of this engine (one engine per GPU / rank) */
	int32_t ctas_per_sm;   /* 0 = let the engine choose from the occupancy query */
	int32_t async_params;  /* 0: cnf2_set_params returns when its copies are done (caller buffers reusable at once);
	                          1: copies from page-locked buffers stay in flight -- keep the arrays unchanged until the
	                          chromosome's cnf2_get_stats* or cnf2_sync returns */
	int32_t in_flight;     /* E-steps kept in flight (their statistics stay readable): 0 = default (2), up to 4 */
	int32_t segment;       /* markers per segment of the sweeps' scratch: the forward pass keeps alpha at segment starts only
	                          and the backward pass recomputes a segment before it walks it, so the scratch does not grow with
	                          the chromosome; 0 = the engine chooses (the whole chromosome while that fits its budget) */
} cnf2_settings;

/* Fills *s with the reference's HEAD settings (settings.h:3-47) and device 0. */
void cnf2_default_settings(cnf2_settings* s);

int cnf2_create(cnf2_engine** out, const cnf2_settings* settings);
void cnf2_destroy(cnf2_engine* e);
const char* cnf2_last_error(const cnf2_engine* e); /* e may be NULL: error of the last failed cnf2_create */

/* Map: marker positions in cM, chromosome boundaries, recombination rates.
 * Replaces the globals markerposes / chromstarts / genrec (cnF2freq.cpp:233-239, set at :7927-7943 and by
 * readalphamap :6669-6685).  chromstarts has n_chrom+1 entries; genrec[gen] is the Haldane exponent
 * (-1/50 at HEAD) used as r = 0.5*(1-exp(genrec[gen]*d)) (:2286, getactrec :777-780). */
int cnf2_set_map(cnf2_engine* e, const double* markerpos_cM, int32_t n_markers, const uint32_t* chromstarts,
                 int32_t n_chrom, const double genrec[3]);

/* Pedigree: replaces individer[] / dous (cnF2freq.cpp:2447-2450) and the individ fields pars, empty,
 * founder, gen, descendants (:853-872).  parents is [n_ind][2] (0 = none).  dous lists the analysed
 * individuals THIS engine works on (a rank's shard of the reference's dous); loglik rows follow its order.
 * The focus-tree tables of fixtrees (:3099-3187) are built here.  fixtrees' sticky side effect
 * (ind->founder = true for analysed individuals without any non-empty ancestor, :3170-3177) is applied
 * for every entry of all_dous (the union over all ranks; pass dous again when not sharded) before any
 * tree is used, i.e. the state the reference is in from its second EM iteration on. */
int cnf2_set_pedigree(cnf2_engine* e, int32_t n_ind, const int32_t* parents, const uint8_t* empty,
                      const uint8_t* founder, const int32_t* gen, const int32_t* descendants, const int32_t* dous,
                      int32_t n_dous, const int32_t* all_dous, int32_t n_all_dous);

/* Parameters of one chromosome for the coming E-step: markerdata (allele pair, values 0 = unknown, 1, 2,
 * 9 = sex/doubled-haploid sentinel, cnF2freq.cpp:225-226, :876), markersure (:878) and haploweight (:887).
 * Arrays are [n_ind][n_markers] (x2 for the pairs); only the columns of `chrom` and the rows of the
 * individuals this engine's focus trees touch are copied (host -> device, asynchronously when the
 * buffers are pinned).  When the parameters of all chromosomes fit in a small part of HBM they stay resident
 * (cnf2_estep may then be called for any chromosome set so far); otherwise only the last chromosome is kept. */
int cnf2_set_params(cnf2_engine* e, int32_t chrom, const uint8_t* alleles, const double* sure,
                    const double* haploweight);

/* The same for arrays already in the engine's row order (cnf2_row_ids) and holding only this chromosome:
 * alleles [n_rows][Mc][2], sure [n_rows][Mc][2], haploweight [n_rows][Mc] -- three unstrided host -> device
 * copies.  The reference keeps these per individual (:876-887), so a binding gathers them either way. */
int cnf2_set_params_packed(cnf2_engine* e, int32_t chrom, const uint8_t* alleles, const double* sure,
                           const double* haploweight);

/* The E-step of one chromosome for all of this engine's analysed individuals: cnF2freq.cpp:5294-5907
 * without the turn scoring (:5668-5871, row f2 of SURVEY.md section 8).  Asynchronous; cnf2_get_stats /
 * cnf2_sync wait for it.  Two E-steps (cnf2_settings.in_flight) are kept in flight (each with its own stream, scratch and statistics
 * block), so the tail of one chromosome's launch overlaps the next chromosome's: call cnf2_estep for chromosome
 * c + 1 before cnf2_get_stats for chromosome c to use that. */
int cnf2_estep(cnf2_engine* e, int32_t chrom);

/* Optional exchange step between ranks (the stale MPI reduce at cnF2freq.cpp:6241-6256): called by
 * cnf2_estep after the kernels are enqueued, with the device buffer holding the statistics rows of the
 * individuals listed in cnf2_set_shared_rows (9 doubles per (row, marker): haplobase, haplocount,
 * infprobs[0][1], [0][2], [1][1], [1][2], homozyg[0], homozyg[1], negshift).  The hook must enqueue a sum-all-reduce
 * over `count` doubles in place on `cuda_stream` (e.g. ncclAllReduce / torch.distributed.all_reduce).  The
 * engine passes one of two streams.  The statistics stream carries every E-step's all-reduce, in cnf2_estep call order.
 * The control stream (cnf2_control_stream) carries the small sums of the block behind a chromosome -- hitnnn of
 * cnf2_mstep_pass (one double), the phase-inversion minima of cnf2_em_pass (two doubles per individual) -- in call order
 * as well.  The two are independent on purpose: behind the statistics of the E-steps issued ahead, a pass's one-double sum
 * (and the host, in cnf2_em_pass) would wait for those E-steps to finish.  A binding therefore uses ONE COMMUNICATOR PER
 * STREAM (a second ncclComm_t / torch.distributed.new_group for the control stream); with a single communicator the
 * collectives of the two streams must be serialised by the binding (torch.distributed does that by itself), which is
 * correct and only gives the overlap away.  Every rank calls the engine in the same order, so each communicator sees its
 * collectives in the same order on every rank. */
typedef int (*cnf2_allreduce_fn)(void* ctx, double* device_buf, size_t count, void* cuda_stream);
int cnf2_set_allreduce(cnf2_engine* e, cnf2_allreduce_fn fn, void* ctx);
/* The control stream the hook is called with for the M-step passes' collectives (a cudaStream_t); see above. */
void* cnf2_control_stream(cnf2_engine* e);
/* Individuals whose statistics rows receive contributions from more than one rank (e.g. the founders of
 * an F2 cross).  Must be the same list, in the same order, on every rank.  Call before cnf2_set_pedigree. */
int cnf2_set_shared_rows(cnf2_engine* e, const int32_t* ids, int32_t n_ids);

/* Statistics of a chromosome among the last two passed to cnf2_estep, copied (assigned, device -> host) into caller
 * arrays shaped like the reference's per-individual vectors: haplobase/haplocount [n_ind][n_markers] (:901-902),
 * infprobs [n_ind][n_markers][2 sides][2 alleles {1,2}] (:895), homozyg [n_ind][n_markers][2] (:897);
 * loglik [n_dous][n_chrom][2] = (max_s L_s, logsumexp_s L_s), the pair printed at :5399-5401.  Only the columns of that chromosome and the rows this engine touched are written.
 * Any pointer may be NULL to skip that output. */
int cnf2_get_stats(cnf2_engine* e, int32_t chrom, double* haplobase, double* haplocount, double* infprobs,
                   double* homozyg, double* loglik);

/* The same statistics in the engine's own row order, one contiguous block per quantity (device -> host copies
 * without a stride: the form for page-locked buffers on a hot loop).  The reference keeps these statistics in
 * per-individual vectors (cnF2freq.cpp:895-902), so a binding scatters rows into individuals either way:
 * cnf2_row_ids gives the individual id of every row (cnf2_row_count of them: the individuals this engine's focus
 * trees touch, shared rows first).  haplobase/haplocount [n_rows][Mc], infprobs [n_rows][Mc][2][2],
 * homozyg [n_rows][Mc][2] with Mc = markers of `chrom`; loglik [n_dous][2] for this chromosome.  Any pointer may
 * be NULL. */
int cnf2_row_count(cnf2_engine* e, int32_t* n_rows);
int cnf2_row_ids(cnf2_engine* e, int32_t* ids);
int cnf2_get_stats_packed(cnf2_engine* e, int32_t chrom, double* haplobase, double* haplocount, double* infprobs,
                          double* homozyg, double* loglik);

/* Phase-inversion scores of the chromosome's E-step (row f2 of SURVEY.md section 8): what the turn scoring
 * (cnF2freq.cpp:5668-5757, aroundturner :498-554; the DOTOULBAR 0 path: the 8 turns of at most one individual) and
 * updatenegshifts (:3640-3715) add to individ::negshift (:891) -- for every member of an analysed individual's focus
 * tree log(sum_s rawvals[no turn][s]) - log(sum_s rawvals[turn of that member][s]), divided by the parent's number of
 * analysed children for grandparents.  cnf2_estep always computes them (HAPLOTYPING && !early && !full, :5668);
 * shared rows are summed over ranks with the other statistics.  [n_ind][n_markers] (columns of `chrom`, rows this
 * engine touches) or, packed, [n_rows][Mc] in cnf2_row_ids order.  The reference zeroes negshift after each
 * iteration's dump (:8187); here every E-step starts from zero. */
int cnf2_get_negshift(cnf2_engine* e, int32_t chrom, double* negshift);
int cnf2_get_negshift_packed(cnf2_engine* e, int32_t chrom, double* negshift);

/* ---- M-step on the device ------------------------------------------------
 * cnf2_mstep makes the pass of cnF2freq.cpp:6356-6367 that the reference runs after chromosome `chrom`'s E-step:
 * processinfprobs (:4179-4323) on the chromosome's fresh statistics -> new markerdata / markersure, then
 * updatehaploweights (:4533-4734, with relskewhmm :4325-4467 and the step solver cappedgd :4040-4177) -> new
 * haploweight, for this chromosome and -- as the reference's block sits inside its chromosome loop and walks all
 * markers -- again for the chromosomes whose pass came earlier in the iteration, on the haplobase / haplocount that
 * pass left behind.  The parameters are updated in place on the device (the next cnf2_estep of the chromosome uses
 * them without an upload); cnf2_get_params reads them back ([n_ind][n_markers] arrays, columns of `chrom`).
 * scalefactor / entropyfactor are the reference's globals (:3573-3574); *hitnnn receives the pass's count of capped
 * steps (:4022-4032), from which the caller runs the step-size control (:6373-6390); with several ranks every rank
 * updates its copy of the shared rows (the all-reduced statistics are identical everywhere) and *hitnnn_shared (may be
 * NULL) is the part of *hitnnn those rows contributed: global hitnnn = sum over ranks of (hitnnn - hitnnn_shared) +
 * hitnnn_shared.  Phase inversions (negshifter, :6336-6340) are applied with cnf2_apply_inversions before the
 * pass; a pass that follows inversions runs with scalefactor 0 (:6341), where lastinved (:6290, the breakathalf
 * argument of cappedgd) has no effect.
 * cnf2_set_mstep_inputs: relhaplo (:889), prior markerdata / markersure (:880-881) as [n_ind][n_markers](x2) arrays,
 * and per individual hasprior (priormarkerdata.size() > 0), children (:5246-5259), empty, descendants.  Needs the
 * parameters of all chromosomes resident.  cnf2_begin_iteration forgets the carried haplobase / haplocount
 * (doit zeroes them, :5222-5234). */
int cnf2_set_mstep_inputs(cnf2_engine* e, const double* relhaplo, const uint8_t* prioralleles, const double* priorsure,
                          const uint8_t* hasprior, const int32_t* children, const uint8_t* empty,
                          const int32_t* descendants);
/* The same inputs in two parts, for hosts that keep their arrays in the engine's row order (cnf2_row_ids): the per-individual
 * attributes ([n_ind] arrays), and per chromosome relhaplo [n_rows][Mc], prior markerdata [n_rows][Mc][2], prior markersure
 * [n_rows][Mc][2] as three unstrided host -> device copies (page-locked buffers with cnf2_settings.async_params stay in
 * flight like cnf2_set_params_packed's). */
int cnf2_set_mstep_rows(cnf2_engine* e, const uint8_t* hasprior, const int32_t* children, const uint8_t* empty,
                        const int32_t* descendants);
int cnf2_set_mstep_chrom_packed(cnf2_engine* e, int32_t chrom, const double* relhaplo, const uint8_t* prioralleles,
                                const double* priorsure);
int cnf2_begin_iteration(cnf2_engine* e);
int cnf2_mstep(cnf2_engine* e, int32_t chrom, double scalefactor, double entropyfactor, int64_t* hitnnn,
               int64_t* hitnnn_shared);
int cnf2_get_params(cnf2_engine* e, int32_t chrom, uint8_t* alleles, double* sure, double* haploweight);
/* The chromosome's parameters in the engine's row order: alleles [n_rows][Mc][2], sure [n_rows][Mc][2], haploweight
 * [n_rows][Mc] -- what the per-iteration dump (cnF2freq.cpp:8157-8192) prints, as three unstrided device -> host copies. */
int cnf2_get_params_packed(cnf2_engine* e, int32_t chrom, uint8_t* alleles, double* sure, double* haploweight);
/* The same without waiting: the copies (into page-locked caller buffers; any pointer may be NULL) are enqueued behind the last
 * cnf2_mstep_pass / cnf2_em_pass / cnf2_apply_inversions issued so far and run beside whatever is issued later; cnf2_sync
 * waits for them.  What a host that writes the per-iteration dump (:8157-8192) uses to take a chromosome's alleles and
 * markersure as soon as its own pass is through -- processinfprobs writes them in that pass only (:6356-6367), later passes
 * and inversions of the iteration touch haploweight alone (:3437-3459, :4533-4734) -- and the haploweights at the
 * iteration's end, so that most of the read-back hides behind the remaining chromosomes' kernels. */
int cnf2_get_params_packed_async(cnf2_engine* e, int32_t chrom, uint8_t* alleles, double* sure, double* haploweight);

/* ---- the EM loop without host round trips (rows a6 / f1 / f2) ---------------------------------------------------
 * The reference's step-size control (cnF2freq.cpp:6333-6392: scalefactor /= 1.1 on a bad pass, *= 1.21 on a good one,
 * *= 0.997 always, from hitnnn and its two-pass history oldhitnnn / oldhitnnn2, :4004-4005) lives on the device:
 * cnf2_set_em_state uploads the globals (scalefactor :3573 = 0.013, entropyfactor :3574 = 1, history 0 at start),
 * cnf2_mstep_pass enqueues -- asynchronously, behind chromosome `chrom`'s E-step -- the same pass as cnf2_mstep with
 * the device's scalefactor, sums hitnnn over ranks (through the cnf2_set_allreduce hook) and runs the control;
 * cnf2_get_em_state waits and reads the state back (hitnnn_last: the global count of the last pass, the number
 * printed at :6392).  any_inversion != 0: the pass of an iteration in which phase inversions were applied -- it
 * runs with scalefactor 0 and leaves scalefactor and the history as they were (:6339-6341, :6385).
 *
 * Phase inversions (negshifter :3437-3459; candidates :6270-6327): cnf2_negshift_minima gives, per row of
 * cnf2_row_ids, the most negative phase-inversion score of the chromosome below -1e-10 and the first marker holding
 * it (chromosome-relative; = Mc when there is none), i.e. minval / minstart of :6272-6284.  Which candidates to keep
 * is host logic (inferiorrelated :3415-3435 over related individuals and libc rand(), :6319); cnf2_apply_inversions
 * applies the kept ones to the device-resident state: haploweight -> 1 - haploweight and haplobase -> haplocount -
 * haplobase from the marker after minstart to the chromosome's end (ids this engine does not hold are skipped).
 * Call it before the cnf2_mstep_pass of the same pass, as the reference does (:6336-6340). */
int cnf2_set_em_state(cnf2_engine* e, double scalefactor, double entropyfactor, int32_t oldhitnnn, int32_t oldhitnnn2);
int cnf2_get_em_state(cnf2_engine* e, double* scalefactor, double* entropyfactor, int32_t* oldhitnnn, int32_t* oldhitnnn2,
                      int64_t* hitnnn_last);
int cnf2_mstep_pass(cnf2_engine* e, int32_t chrom, int32_t any_inversion);
int cnf2_negshift_minima(cnf2_engine* e, int32_t chrom, double* minval, int32_t* minstart);
int cnf2_apply_inversions(cnf2_engine* e, int32_t chrom, const int32_t* ids, const int32_t* minstart, int32_t n);

/* The whole block the reference runs after a chromosome's E-step (cnF2freq.cpp:6239-6392) in one call: phase-inversion
 * candidates of every chromosome analysed so far in this iteration (minima exchanged between ranks through the
 * cnf2_set_allreduce hook; inferiorrelated / arerelated :916-946 and rand() on the host, in the reference's order:
 * individuals by id, chromosomes inside), negshifter on all of them, then cnf2_mstep_pass.  The applied inversions
 * are returned for the "Inv: <id> <marker>" lines (:3450; marker = the first inverted one, over all chromosomes; at
 * most cap entries are written, *n_inv is the full count).  Uses libc rand() exactly as the reference does (:6319),
 * so a host that never seeds or calls rand() elsewhere reproduces the reference binary's decisions.
 * cnf2_begin_iteration forgets the candidates (negshiftcands is local to doit, :5261). */
int cnf2_em_pass(cnf2_engine* e, int32_t chrom, int32_t* inv_ids, int32_t* inv_markers, int32_t cap, int32_t* n_inv);

/* ---- pre-processing (row f3): postmarkerdata, cnF2freq.cpp:3191-3412 ----------------------------------------------------
 * What the reference runs once between reading its inputs and the first EM iteration (main, :8083-8085, with
 * CORRECTIONINFERENCE true): genotype inference for untyped relatives to a fixed point -- fixkid :1469-1487, fixparents
 * :1392-1467 and the corrections drawn from the evidence it collects :3283-3351 --, the descendant and child counts
 * (:3228-3277), addvariance :1489-1558 and lockhaplos :3045-3079 (the phase of every individual is locked at the marker of
 * largest variance of each chromosome).  The two emission scans (fixparents: up to 64 x 128 emissions per individual-marker;
 * addvariance: 2 x 2 x 2 x 64 x 64 x 2) run on the device for all individuals at once; the bookkeeping between them follows the
 * reference's single-threaded order.  Needs cnf2_set_map.  All arrays are [n_ind](...) / [n_ind][n_markers](x2), caller
 * owned: alleles / sure / haploweight in-out (as the readers left them -> as the first EM iteration finds them),
 * descendants in-out (zero on entry for a fresh run), children and founder (fixtrees' sticky flag, :3170-3177: pass it to
 * cnf2_set_pedigree) out.  exists: the individual exists (individer[] non-null).  log (may be NULL) receives the
 * reference's progress lines ("Correction at", "Fixing point", to stdout = 0; "Number of corrected genotypes", "Clearing"
 * to stderr = 1) in its order. */
typedef void (*cnf2_log_fn)(void* ctx, int32_t stream, const char* line);
int cnf2_preprocess(cnf2_engine* e, int32_t n_ind, const int32_t* parents, const uint8_t* exists, const uint8_t* empty,
                    uint8_t* alleles, double* sure, double* haploweight, int32_t* descendants, int32_t* children,
                    uint8_t* founder, cnf2_log_fn log, void* log_ctx);
/* addvariance's 65,536 emission walks per (individual, marker) feed one decision only: which marker of a chromosome holds the
 * largest variance (lockhaplos :3058, a strict `>`).  With SELFING false cnf2_preprocess therefore ranks the markers with the
 * factorised form of those sums (~500 node evaluations per cell, equal up to rounding) and walks the literal sums, in the
 * reference's order of operations, only for the markers that can hold the maximum: those within 1e-9 of the largest ranked
 * value, one per distinct set of genotypes in the focus tree -- the decision is the literal one.  This returns how many
 * (individual, marker) cells the last cnf2_preprocess evaluated literally (all of them under SELFING or with
 * CNF2_PRE_LITERAL=1 in the environment); -1 for a NULL engine.  Replaces nothing in the reference (a diagnostic). */
int64_t cnf2_preprocess_literal_cells(const cnf2_engine* e);

/* Block until everything enqueued by this engine has finished; returns the first asynchronous error. */
int cnf2_sync(cnf2_engine* e);

/* Measurement hooks (bench.py): device time of the last cnf2_estep's kernels in milliseconds (CUDA events
 * on the engine's stream), number of kernel launches it made, and the engine's stream handle. */
int cnf2_last_estep_ms(cnf2_engine* e, float* ms);
int cnf2_last_estep_launches(cnf2_engine* e, int32_t* n);
/* Device time of the E-step kernels summed over the cnf2_estep calls since the last reset (CUDA events on each launch's
 * own stream, so a launch that shares the GPU with other launches or with an M-step pass counts its full duration), their
 * number, and the number of kernels of any kind this engine launched.  Waits for the launches in flight. */
int cnf2_estep_times(cnf2_engine* e, double* sum_ms, int64_t* n_launches, int64_t* n_kernels, int32_t reset);
/* Counters of the step solver (cappedgd, cnF2freq.cpp:4040-4177) since the last reset, four values: gradient evaluations
 * (one at the start, one per bisection midpoint, 15 per Gauss-Legendre integral) of [0] processinfprobs and [1]
 * updatehaploweights; solves that ended on integral-free steps and were [2] confirmed by their closing integral, [3] sent
 * back to repeat those steps with their integrals (see ms_cappedgd_warp).  Waits for the passes in flight. */
int cnf2_mstep_counters(cnf2_engine* e, uint64_t* evals, int32_t reset);
void* cnf2_stream(cnf2_engine* e);
/* CUDA events on the engine's stream: cnf2_mark records event `slot` (0..3); cnf2_elapsed_ms waits for
 * event `to` and returns the device time between two recorded events. */
int cnf2_mark(cnf2_engine* e, int32_t slot);
int cnf2_elapsed_ms(cnf2_engine* e, int32_t from, int32_t to, float* ms);

/* Microbenchmarks that define the roofline denominators on this box:
 * register-resident DFMA throughput in TFLOP/s and device-to-device copy bandwidth in GB/s. */
int cnf2_measure_peaks(cnf2_engine* e, double* dfma_tflops, double* copy_gbs);

/* ---- an entry point that allocates nothing ------------------------------------------------------------------------------
 * No counterpart in the reference: added for callers that cannot call cudaMalloc
 * (code under CUDA graph capture, tight serving loops, harnesses that forbid allocation).  The minima of
 * cnf2_negshift_minima -- per row the most negative phase-inversion score below -1e-10 and the first marker holding it,
 * minstart = Mc where there is none -- computed from and into the CALLER's device buffers on the caller's stream, on the
 * current device: negshift_dev [n_rows][Mc] doubles (the layout of cnf2_get_negshift_packed), minval_dev [n_rows] doubles,
 * minstart_dev [n_rows] int32, stream a cudaStream_t (NULL = the default stream).  No engine handle is needed.  The call
 * makes no cudaMalloc / cudaFree and returns without waiting for the kernels.  Their scratch (one partial minimum per row and
 * chunk of 256 markers) is a static 8 MiB device array inside the library, held by every GPU that loads the library until the
 * process exits: n_rows * ceil(Mc / 256) may not exceed 1,048,576 (CNF2_ERR_BAD_ARG beyond that; the message is returned by
 * cnf2_last_error(NULL)).  All calls share that array: calls that overlap on the GPU (different streams without a
 * dependency between them) race on it; calls on one stream, or ordered by events, are safe.  Nothing in it has to start at
 * zero -- every cell a call reads it has written first.  The first call on each device looks the array's address up
 * (cudaGetSymbolAddress); make that call outside a graph capture. */
int cnf2_negshift_minima_noalloc(const double* negshift_dev, int32_t n_rows, int32_t Mc, double* minval_dev, int32_t* minstart_dev,
                                 void* stream);

#ifdef __cplusplus
}
#endif
#endif
