"""Host-side binding of the C ABI in include/cnf2_engine.h (ctypes; no torch types cross the boundary).

This is the Python mirror of what a C++ host does around the cut in doit() (cnF2freq.cpp:5283-5907):
per chromosome, hand the engine the current parameters, run the E-step, take the EM sufficient
statistics back.  There is no CPU path: if the CUDA library is missing or no GPU is present the calls
raise.
"""
import ctypes
import os

import numpy as np

_PKG = os.path.dirname(os.path.abspath(__file__))
# CNF2_ENGINE_LIB: another build of the same library (A/B experiments, instrumented builds); never a CPU path
LIB_PATH = os.environ.get("CNF2_ENGINE_LIB") or os.path.join(_PKG, "lib", "libcnf2_engine.so")

SYMBOLS = [
    "cnf2_default_settings", "cnf2_create", "cnf2_destroy", "cnf2_last_error", "cnf2_set_map", "cnf2_set_pedigree",
    "cnf2_set_params", "cnf2_estep", "cnf2_set_allreduce", "cnf2_set_shared_rows", "cnf2_get_stats", "cnf2_sync",
    "cnf2_last_estep_ms", "cnf2_last_estep_launches", "cnf2_stream", "cnf2_measure_peaks", "cnf2_mark",
    "cnf2_elapsed_ms", "cnf2_row_count", "cnf2_row_ids", "cnf2_get_stats_packed", "cnf2_set_params_packed", "cnf2_set_mstep_inputs",
    "cnf2_begin_iteration", "cnf2_mstep", "cnf2_get_params", "cnf2_get_negshift", "cnf2_get_negshift_packed",
    "cnf2_set_em_state", "cnf2_get_em_state", "cnf2_mstep_pass", "cnf2_negshift_minima", "cnf2_apply_inversions", "cnf2_em_pass",
    "cnf2_set_mstep_rows", "cnf2_set_mstep_chrom_packed", "cnf2_get_params_packed", "cnf2_estep_times", "cnf2_mstep_counters",
    "cnf2_preprocess", "cnf2_negshift_minima_noalloc", "cnf2_preprocess_literal_cells", "cnf2_control_stream",
    "cnf2_get_params_packed_async",
]


class Settings(ctypes.Structure):
    _fields_ = [(n, ctypes.c_int32) for n in
                ("numgen", "selfing", "relskewstates", "haplotyping", "doinfprobs", "device", "ctas_per_sm", "async_params", "in_flight", "segment")]


ALLREDUCE_FN = ctypes.CFUNCTYPE(ctypes.c_int, ctypes.c_void_p, ctypes.c_void_p, ctypes.c_size_t, ctypes.c_void_p)
LOG_FN = ctypes.CFUNCTYPE(None, ctypes.c_void_p, ctypes.c_int32, ctypes.c_char_p)


class EngineError(RuntimeError):
    pass


_lib = None


def load_library():
    """Loads libcnf2_engine.so; raises (never falls back) when it has not been built."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise EngineError(f"{LIB_PATH} is missing: run `python -c 'import __graft_entry__ as g; g.build()'` "
                              "(the engine has no CPU fallback)")
        lib = ctypes.CDLL(LIB_PATH)
        lib.cnf2_last_error.restype = ctypes.c_char_p
        lib.cnf2_last_error.argtypes = [ctypes.c_void_p]
        lib.cnf2_stream.restype = ctypes.c_void_p
        lib.cnf2_stream.argtypes = [ctypes.c_void_p]
        lib.cnf2_control_stream.restype = ctypes.c_void_p
        lib.cnf2_control_stream.argtypes = [ctypes.c_void_p]
        lib.cnf2_destroy.restype = None
        lib.cnf2_destroy.argtypes = [ctypes.c_void_p]
        lib.cnf2_default_settings.restype = None
        _lib = lib
    return _lib


def negshift_minima_noalloc(negshift_dev, n_rows, mc, minval_dev, minstart_dev, stream=0):
    """cnf2_negshift_minima_noalloc: device addresses (ints) of the caller's buffers, negshift [n_rows][mc] f64 in,
    minval [n_rows] f64 and minstart [n_rows] i32 out, on `stream` (a cudaStream_t as int) of the current device.  Allocates
    nothing; returns once the two kernels are enqueued."""
    lib = load_library()
    rc = lib.cnf2_negshift_minima_noalloc(ctypes.c_void_p(negshift_dev), ctypes.c_int32(n_rows), ctypes.c_int32(mc),
                                          ctypes.c_void_p(minval_dev), ctypes.c_void_p(minstart_dev), ctypes.c_void_p(stream))
    if rc != 0:
        raise EngineError(f"cnf2 error {rc}: {lib.cnf2_last_error(None).decode()}")


def _ptr(a):
    return None if a is None else a.ctypes.data_as(ctypes.c_void_p)


def _buf(a, dtype, size, what):
    """A caller-owned array the library reads or fills through a raw pointer: dtype, C-contiguity and element count are
    checked here, because the C side cannot."""
    if a is None:
        return None
    if not isinstance(a, np.ndarray) or a.dtype != dtype or not a.flags.c_contiguous or a.size != size:
        raise EngineError(f"{what}: needs a C-contiguous numpy array of {np.dtype(dtype).name} with {size} elements, got "
                          f"{type(a).__name__}" + (f" {a.dtype} size {a.size} contiguous={a.flags.c_contiguous}" if isinstance(a, np.ndarray) else ""))
    return a.ctypes.data_as(ctypes.c_void_p)


class Engine:
    """One engine = one GPU (one rank).  Method names follow the C ABI."""

    def __init__(self, device=0, ctas_per_sm=0, async_params=False, in_flight=0, segment=0, selfing=False):
        """selfing: the reference's compile-time knob SELFING (settings.h:14) -- 256 states, PlantImpute lines listed with a
        generation above 2 (demo.sh:18-20); every unit then runs through the literal kernel."""
        self.lib = load_library()
        s = Settings()
        self.lib.cnf2_default_settings(ctypes.byref(s))
        s.selfing = 1 if selfing else 0
        s.device = device
        s.ctas_per_sm = ctas_per_sm
        s.async_params = 1 if async_params else 0
        s.in_flight = in_flight
        s.segment = segment
        self.h = ctypes.c_void_p()
        rc = self.lib.cnf2_create(ctypes.byref(self.h), ctypes.byref(s))
        if rc != 0:
            raise EngineError(f"cnf2_create failed ({rc}): {self.lib.cnf2_last_error(None).decode()}")
        self._keep = []
        self.n_markers = self.n_chrom = self.n_ind = self.n_dous = 0
        self.chromstarts = None

    def _check(self, rc):
        if rc != 0:
            raise EngineError(f"cnf2 error {rc}: {self.lib.cnf2_last_error(self.h).decode()}")

    def close(self):
        if self.h:
            self.lib.cnf2_destroy(self.h)
            self.h = ctypes.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def set_map(self, markerpos, chromstarts, genrec):
        markerpos = np.ascontiguousarray(markerpos, np.float64)
        chromstarts = np.ascontiguousarray(chromstarts, np.uint32)
        genrec = np.ascontiguousarray(genrec, np.float64)
        self.n_markers, self.n_chrom = len(markerpos), len(chromstarts) - 1
        self.chromstarts = chromstarts.copy()
        self._check(self.lib.cnf2_set_map(self.h, _ptr(markerpos), ctypes.c_int32(self.n_markers), _ptr(chromstarts),
                                          ctypes.c_int32(self.n_chrom), _ptr(genrec)))

    def set_shared_rows(self, ids):
        ids = np.ascontiguousarray(ids, np.int32)
        self._check(self.lib.cnf2_set_shared_rows(self.h, _ptr(ids), ctypes.c_int32(len(ids))))

    def set_allreduce(self, fn):
        """fn(device_ptr:int, count:int, cuda_stream:int) -> None; must enqueue an in-place fp64 sum all-reduce."""
        def tramp(ctx, buf, count, stream):
            try:
                fn(int(buf), int(count), int(stream or 0))
                return 0
            except Exception as ex:  # surfaced as CNF2_ERR_CUDA by the engine
                print("cnf2 all-reduce hook failed:", ex)
                return 1
        cb = ALLREDUCE_FN(tramp)
        self._keep.append(cb)
        self._check(self.lib.cnf2_set_allreduce(self.h, cb, None))

    def control_stream(self):
        """The cudaStream_t (int) the hook is called with for the M-step passes' small collectives (cnf2_control_stream): a
        binding gives it a communicator of its own."""
        return int(self.lib.cnf2_control_stream(self.h) or 0)

    def set_pedigree(self, parents, empty, founder, gen, descendants, dous, all_dous=None):
        parents = np.ascontiguousarray(parents, np.int32)
        empty = np.ascontiguousarray(empty, np.uint8)
        founder = np.ascontiguousarray(founder, np.uint8)
        gen = np.ascontiguousarray(gen, np.int32)
        descendants = np.ascontiguousarray(descendants, np.int32)
        dous = np.ascontiguousarray(dous, np.int32)
        all_dous = dous if all_dous is None else np.ascontiguousarray(all_dous, np.int32)
        self.n_ind, self.n_dous = parents.shape[0], len(dous)
        self._check(self.lib.cnf2_set_pedigree(
            self.h, ctypes.c_int32(self.n_ind), _ptr(parents), _ptr(empty), _ptr(founder), _ptr(gen), _ptr(descendants),
            _ptr(dous), ctypes.c_int32(len(dous)), _ptr(all_dous), ctypes.c_int32(len(all_dous))))

    def set_params(self, chrom, alleles, sure, haploweight):
        """alleles u8 [n_ind][M][2], sure f64 [n_ind][M][2], haploweight f64 [n_ind][M] (numpy, C-contiguous)."""
        for a, dt, n in ((alleles, np.uint8, 2), (sure, np.float64, 2), (haploweight, np.float64, 1)):
            if a.dtype != dt or not a.flags.c_contiguous or a.size != self.n_ind * self.n_markers * n:
                raise EngineError("set_params: arrays must be C-contiguous [n_ind][n_markers](x2) of the documented dtype")
        self._check(self.lib.cnf2_set_params(self.h, ctypes.c_int32(chrom), _ptr(alleles), _ptr(sure), _ptr(haploweight)))

    def set_params_packed(self, chrom, alleles, sure, haploweight):
        """Row-ordered (row_ids()) arrays of this chromosome only: alleles u8 [n_rows][Mc][2], sure f64 [n_rows][Mc][2],
        haploweight f64 [n_rows][Mc]."""
        n = self.n_rows() * self.mc(chrom)
        self._check(self.lib.cnf2_set_params_packed(self.h, ctypes.c_int32(chrom), _buf(alleles, np.uint8, 2 * n, "set_params_packed alleles"),
                                                    _buf(sure, np.float64, 2 * n, "set_params_packed sure"),
                                                    _buf(haploweight, np.float64, n, "set_params_packed haploweight")))

    def estep(self, chrom):
        self._check(self.lib.cnf2_estep(self.h, ctypes.c_int32(chrom)))

    def mc(self, chrom):
        return int(self.chromstarts[chrom + 1]) - int(self.chromstarts[chrom])

    def n_rows(self):
        n = ctypes.c_int32()
        self._check(self.lib.cnf2_row_count(self.h, ctypes.byref(n)))
        return int(n.value)

    def get_stats(self, chrom, haplobase=None, haplocount=None, infprobs=None, homozyg=None, loglik=None):
        km = self.n_ind * self.n_markers
        self._check(self.lib.cnf2_get_stats(self.h, ctypes.c_int32(chrom), _buf(haplobase, np.float64, km, "get_stats haplobase"),
                                            _buf(haplocount, np.float64, km, "get_stats haplocount"),
                                            _buf(infprobs, np.float64, 4 * km, "get_stats infprobs"),
                                            _buf(homozyg, np.float64, 2 * km, "get_stats homozyg"),
                                            _buf(loglik, np.float64, 2 * self.n_dous * self.n_chrom, "get_stats loglik")))

    def get_negshift(self, chrom, negshift):
        """Phase-inversion scores of the chromosome's E-step into negshift f64 [n_ind][n_markers] (columns of `chrom`)."""
        self._check(self.lib.cnf2_get_negshift(self.h, ctypes.c_int32(chrom),
                                               _buf(negshift, np.float64, self.n_ind * self.n_markers, "get_negshift")))

    def get_negshift_packed(self, chrom, negshift):
        self._check(self.lib.cnf2_get_negshift_packed(self.h, ctypes.c_int32(chrom),
                                                      _buf(negshift, np.float64, self.n_rows() * self.mc(chrom), "get_negshift_packed")))

    def row_ids(self):
        """Individual id of every statistics row of this engine (shared rows first)."""
        n = self.n_rows()
        ids = np.zeros(n, np.int32)
        if n:
            self._check(self.lib.cnf2_row_ids(self.h, _ptr(ids)))
        return ids

    def get_stats_packed(self, chrom, haplobase=None, haplocount=None, infprobs=None, homozyg=None, loglik=None):
        """Row-ordered contiguous blocks ([n_rows][Mc]...; see include/cnf2_engine.h): unstrided device -> host copies."""
        n = self.n_rows() * self.mc(chrom)
        self._check(self.lib.cnf2_get_stats_packed(self.h, ctypes.c_int32(chrom), _buf(haplobase, np.float64, n, "get_stats_packed haplobase"),
                                                   _buf(haplocount, np.float64, n, "get_stats_packed haplocount"),
                                                   _buf(infprobs, np.float64, 4 * n, "get_stats_packed infprobs"),
                                                   _buf(homozyg, np.float64, 2 * n, "get_stats_packed homozyg"),
                                                   _buf(loglik, np.float64, 2 * self.n_dous, "get_stats_packed loglik")))

    # -- pre-processing (row f3) ----------------------------------------------------------------------
    def preprocess(self, parents, exists, empty, alleles, sure, haploweight, descendants=None, log=None):
        """postmarkerdata (cnF2freq.cpp:3191-3412) on arrays as the readers left them; alleles / sure / haploweight are
        updated in place.  Returns (descendants, children, founder).  log(stream, line) receives the progress lines."""
        c = np.ascontiguousarray
        parents, exists, empty = c(parents, np.int32), c(exists, np.uint8), c(empty, np.uint8)
        k, km = parents.shape[0], parents.shape[0] * self.n_markers
        desc = np.zeros(k, np.int32) if descendants is None else c(descendants, np.int32).copy()
        children, founder = np.zeros(k, np.int32), np.zeros(k, np.uint8)
        cb = LOG_FN((lambda ctx, stream, line: log(int(stream), line.decode())) if log else (lambda ctx, stream, line: None))
        self._check(self.lib.cnf2_preprocess(
            self.h, ctypes.c_int32(k), _ptr(parents), _buf(exists, np.uint8, k, "exists"), _buf(empty, np.uint8, k, "empty"),
            _buf(alleles, np.uint8, 2 * km, "preprocess alleles"), _buf(sure, np.float64, 2 * km, "preprocess sure"),
            _buf(haploweight, np.float64, km, "preprocess haploweight"), _ptr(desc), _ptr(children), _ptr(founder), cb, None))
        return desc, children, founder

    def preprocess_literal_cells(self):
        """(individual, marker) cells whose addvariance sums the last preprocess() walked term by term (cnf2_preprocess_literal_cells)."""
        f = self.lib.cnf2_preprocess_literal_cells
        f.restype = ctypes.c_int64
        return int(f(self.h))

    # -- M-step on the device (rows a6 / f1) ---------------------------------------------------------
    def set_mstep_inputs(self, relhaplo, prioralleles, priorsure, hasprior, children, empty, descendants):
        c = np.ascontiguousarray
        km, k = self.n_ind * self.n_markers, self.n_ind
        self._check(self.lib.cnf2_set_mstep_inputs(
            self.h, _buf(c(relhaplo, np.float64), np.float64, km, "relhaplo"), _buf(c(prioralleles, np.uint8), np.uint8, 2 * km, "prioralleles"),
            _buf(c(priorsure, np.float64), np.float64, 2 * km, "priorsure"), _buf(c(hasprior, np.uint8), np.uint8, k, "hasprior"),
            _buf(c(children, np.int32), np.int32, k, "children"), _buf(c(empty, np.uint8), np.uint8, k, "empty"),
            _buf(c(descendants, np.int32), np.int32, k, "descendants")))

    def set_mstep_rows(self, hasprior, children, empty, descendants):
        c, k = np.ascontiguousarray, self.n_ind
        self._check(self.lib.cnf2_set_mstep_rows(
            self.h, _buf(c(hasprior, np.uint8), np.uint8, k, "hasprior"), _buf(c(children, np.int32), np.int32, k, "children"),
            _buf(c(empty, np.uint8), np.uint8, k, "empty"), _buf(c(descendants, np.int32), np.int32, k, "descendants")))

    def set_mstep_chrom_packed(self, chrom, relhaplo, prioralleles, priorsure):
        """Row-ordered arrays of this chromosome: relhaplo f64 [n_rows][Mc], prioralleles u8 [n_rows][Mc][2], priorsure f64
        [n_rows][Mc][2]."""
        n = self.n_rows() * self.mc(chrom)
        self._check(self.lib.cnf2_set_mstep_chrom_packed(
            self.h, ctypes.c_int32(chrom), _buf(relhaplo, np.float64, n, "set_mstep_chrom_packed relhaplo"),
            _buf(prioralleles, np.uint8, 2 * n, "set_mstep_chrom_packed prioralleles"),
            _buf(priorsure, np.float64, 2 * n, "set_mstep_chrom_packed priorsure")))

    def begin_iteration(self):
        self._check(self.lib.cnf2_begin_iteration(self.h))

    def mstep(self, chrom, scalefactor, entropyfactor, split=False):
        """The reference's M-step pass after chromosome `chrom`'s E-step; returns hitnnn (split=True: (hitnnn, the
        part of it from shared rows) -- see include/cnf2_engine.h for the multi-rank sum)."""
        hit, sh = ctypes.c_int64(), ctypes.c_int64()
        self._check(self.lib.cnf2_mstep(self.h, ctypes.c_int32(chrom), ctypes.c_double(scalefactor),
                                        ctypes.c_double(entropyfactor), ctypes.byref(hit), ctypes.byref(sh)))
        return (int(hit.value), int(sh.value)) if split else int(hit.value)

    # -- the EM loop on the device (rows a6 / f1 / f2): step-size control, phase inversions ------------------
    def set_em_state(self, scalefactor=0.013, entropyfactor=1.0, oldhitnnn=0, oldhitnnn2=0):
        self._check(self.lib.cnf2_set_em_state(self.h, ctypes.c_double(scalefactor), ctypes.c_double(entropyfactor),
                                               ctypes.c_int32(int(oldhitnnn)), ctypes.c_int32(int(oldhitnnn2))))

    def get_em_state(self):
        """(scalefactor, entropyfactor, oldhitnnn, oldhitnnn2, hitnnn of the last pass); waits for the passes in flight."""
        sf, ent = ctypes.c_double(), ctypes.c_double()
        o1, o2, hit = ctypes.c_int32(), ctypes.c_int32(), ctypes.c_int64()
        self._check(self.lib.cnf2_get_em_state(self.h, ctypes.byref(sf), ctypes.byref(ent), ctypes.byref(o1), ctypes.byref(o2),
                                               ctypes.byref(hit)))
        return float(sf.value), float(ent.value), int(o1.value), int(o2.value), int(hit.value)

    def mstep_pass(self, chrom, any_inversion=False):
        self._check(self.lib.cnf2_mstep_pass(self.h, ctypes.c_int32(chrom), ctypes.c_int32(1 if any_inversion else 0)))

    def negshift_minima(self, chrom):
        """(minval f64 [n_rows], minstart i32 [n_rows]) in row_ids() order; minstart = Mc where no score is below -1e-10."""
        n = self.n_rows()
        mv, ms = np.zeros(n, np.float64), np.zeros(n, np.int32)
        self._check(self.lib.cnf2_negshift_minima(self.h, ctypes.c_int32(chrom), _ptr(mv), _ptr(ms)))
        return mv, ms

    def apply_inversions(self, chrom, ids, minstart):
        ids = np.ascontiguousarray(ids, np.int32)
        minstart = np.ascontiguousarray(minstart, np.int32)
        if len(ids) != len(minstart):
            raise EngineError("apply_inversions: ids and minstart differ in length")
        self._check(self.lib.cnf2_apply_inversions(self.h, ctypes.c_int32(chrom), _ptr(ids), _ptr(minstart), ctypes.c_int32(len(ids))))

    def em_pass(self, chrom, cap=4096):
        """The reference's block behind a chromosome's E-step (candidates, inversions, M-step pass, step-size control);
        returns the applied inversions as a list of (individual id, first inverted marker)."""
        ids, mk, n = np.zeros(cap, np.int32), np.zeros(cap, np.int32), ctypes.c_int32()
        self._check(self.lib.cnf2_em_pass(self.h, ctypes.c_int32(chrom), _ptr(ids), _ptr(mk), ctypes.c_int32(cap), ctypes.byref(n)))
        k = min(int(n.value), cap)
        return list(zip(ids[:k].tolist(), mk[:k].tolist()))

    def get_params(self, chrom, alleles, sure, haploweight):
        km = self.n_ind * self.n_markers
        self._check(self.lib.cnf2_get_params(self.h, ctypes.c_int32(chrom), _buf(alleles, np.uint8, 2 * km, "get_params alleles"),
                                             _buf(sure, np.float64, 2 * km, "get_params sure"),
                                             _buf(haploweight, np.float64, km, "get_params haploweight")))

    def get_params_packed(self, chrom, alleles, sure, haploweight):
        n = self.n_rows() * self.mc(chrom)
        self._check(self.lib.cnf2_get_params_packed(self.h, ctypes.c_int32(chrom), _buf(alleles, np.uint8, 2 * n, "get_params_packed alleles"),
                                                    _buf(sure, np.float64, 2 * n, "get_params_packed sure"),
                                                    _buf(haploweight, np.float64, n, "get_params_packed haploweight")))

    def get_params_packed_async(self, chrom, alleles=None, sure=None, haploweight=None):
        """cnf2_get_params_packed_async: enqueues the read-back (page-locked buffers) and returns; sync() waits for it."""
        n = self.n_rows() * self.mc(chrom)
        self._check(self.lib.cnf2_get_params_packed_async(
            self.h, ctypes.c_int32(chrom), _buf(alleles, np.uint8, 2 * n, "get_params_packed_async alleles"),
            _buf(sure, np.float64, 2 * n, "get_params_packed_async sure"), _buf(haploweight, np.float64, n, "get_params_packed_async haploweight")))

    def estep_times(self, reset=False):
        """(summed device ms of the E-step kernels, their launches, kernels of any kind launched) since the last reset."""
        ms, n, k = ctypes.c_double(), ctypes.c_int64(), ctypes.c_int64()
        self._check(self.lib.cnf2_estep_times(self.h, ctypes.byref(ms), ctypes.byref(n), ctypes.byref(k), ctypes.c_int32(1 if reset else 0)))
        return float(ms.value), int(n.value), int(k.value)

    def mstep_counters(self, reset=False):
        """Step-solver counters since the last reset: gradient evaluations of (processinfprobs, updatehaploweights), solves
        ended on integral-free steps (confirmed by the closing integral, sent back)."""
        ev = (ctypes.c_uint64 * 4)()
        self._check(self.lib.cnf2_mstep_counters(self.h, ev, ctypes.c_int32(1 if reset else 0)))
        return int(ev[0]), int(ev[1]), int(ev[2]), int(ev[3])

    def sync(self):
        self._check(self.lib.cnf2_sync(self.h))

    def last_estep_ms(self):
        ms = ctypes.c_float()
        self._check(self.lib.cnf2_last_estep_ms(self.h, ctypes.byref(ms)))
        return float(ms.value)

    def last_estep_launches(self):
        n = ctypes.c_int32()
        self._check(self.lib.cnf2_last_estep_launches(self.h, ctypes.byref(n)))
        return int(n.value)

    def stream(self):
        return int(self.lib.cnf2_stream(self.h) or 0)

    def mark(self, slot):
        self._check(self.lib.cnf2_mark(self.h, ctypes.c_int32(slot)))

    def elapsed_ms(self, a, b):
        ms = ctypes.c_float()
        self._check(self.lib.cnf2_elapsed_ms(self.h, ctypes.c_int32(a), ctypes.c_int32(b), ctypes.byref(ms)))
        return float(ms.value)

    def measure_peaks(self):
        a, b = ctypes.c_double(), ctypes.c_double()
        self._check(self.lib.cnf2_measure_peaks(self.h, ctypes.byref(a), ctypes.byref(b)))
        return float(a.value), float(b.value)

    # -- convenience: a whole E-step over all chromosomes, shaped like the oracle's outputs ------------
    def estep_all(self, alleles, sure, haploweight, out=None):
        K, M, C = self.n_ind, self.n_markers, self.n_chrom
        if out is None:
            out = dict(haplobase=np.zeros((K, M)), haplocount=np.zeros((K, M)), infprobs=np.zeros((K, M, 2, 2)),
                       homozyg=np.zeros((K, M, 2)), loglik=np.zeros((self.n_dous, C, 2)), negshift=np.zeros((K, M)))
        for c in range(C):
            self.set_params(c, alleles, sure, haploweight)
            self.estep(c)
            self.get_stats(c, out["haplobase"], out["haplocount"], out["infprobs"], out["homozyg"], out["loglik"])
            if "negshift" in out:
                self.get_negshift(c, out["negshift"])
        return out
