"""ctypes binding of the CPU oracle (oracle/cnf2_oracle.cpp).  TEST INFRASTRUCTURE ONLY."""
import ctypes
import os
import subprocess

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
_LIB = os.path.join(ROOT, "oracle", "_build", "libcnf2oracle.so")
# the same sources compiled as the SELFING = true instantiation (a compile-time knob of the reference, settings.h:14)
_LIB_SELFING = os.path.join(ROOT, "oracle", "_build", "libcnf2oracle_selfing.so")


def build_oracle(force=False, selfing=False):
    srcs = [os.path.join(ROOT, "oracle", f) for f in ("cnf2_oracle.cpp", "cnf2_mstep_oracle.cpp", "cnf2_preprocess_oracle.cpp")]
    out = _LIB_SELFING if selfing else _LIB
    if force or not os.path.exists(out) or os.path.getmtime(out) < max(os.path.getmtime(s) for s in srcs):
        os.makedirs(os.path.dirname(out), exist_ok=True)
        subprocess.check_call(["g++", "-std=c++17", "-O2", "-fopenmp", "-fPIC", "-shared"] +
                              (["-DCNF2_ORACLE_SELFING=1"] if selfing else []) + ["-o", out] + srcs)
    return out


_lib = None
_lib_selfing = None


def lib(selfing=False):
    global _lib, _lib_selfing
    if selfing:
        if _lib_selfing is None:
            _lib_selfing = ctypes.CDLL(build_oracle(selfing=True))
        return _lib_selfing
    if _lib is None:
        _lib = ctypes.CDLL(build_oracle())
    return _lib


def _p(a):
    return a.ctypes.data_as(ctypes.c_void_p) if a is not None else None


class Problem:
    """Flat-array view of a pedigree + map + parameters (layout of include/cnf2_engine.h)."""

    def __init__(self, par, empty, founder, gen, descendants, dous, markerpos, chromstarts, genrec, alleles, sure,
                 haploweight):
        self.par = np.ascontiguousarray(par, np.int32)
        self.K = self.par.shape[0]
        self.empty = np.ascontiguousarray(empty, np.uint8)
        self.founder = np.ascontiguousarray(founder, np.uint8)
        self.gen = np.ascontiguousarray(gen, np.int32)
        self.descendants = np.ascontiguousarray(descendants, np.int32)
        self.dous = np.ascontiguousarray(dous, np.int32)
        self.markerpos = np.ascontiguousarray(markerpos, np.float64)
        self.M = self.markerpos.shape[0]
        self.chromstarts = np.ascontiguousarray(chromstarts, np.uint32)
        self.C = self.chromstarts.shape[0] - 1
        self.genrec = np.ascontiguousarray(genrec, np.float64)
        self.alleles = np.ascontiguousarray(alleles, np.uint8).reshape(self.K, self.M, 2)
        self.sure = np.ascontiguousarray(sure, np.float64).reshape(self.K, self.M, 2)
        self.haploweight = np.ascontiguousarray(haploweight, np.float64).reshape(self.K, self.M)

    @classmethod
    def from_ref_dump(cls, problem, params, founder=None):
        par = np.stack([problem["par0"], problem["par1"]], axis=1)
        return cls(par, problem["empty"], params["founder"] if founder is None else founder, problem["gen"],
                   problem["descendants"], problem["dous"], problem["markerposes"], problem["chromstarts"],
                   problem["genrec"], params["alleles"], params["sure"], params["haploweight"])


def steady_founder(pb):
    f = pb.founder.copy()
    lib().cnf2_oracle_steady_founder(ctypes.c_int(pb.K), ctypes.c_int(len(pb.dous)), _p(pb.par), _p(pb.empty),
                                     _p(pb.dous), _p(f))
    return f


def doit_children(par, dous):
    """individ::children as doit rebuilds it from the analysed individuals (cnF2freq.cpp:5246-5259)."""
    par = np.asarray(par)
    ch = np.bincount(par[np.asarray(dous, np.int64)].ravel(), minlength=par.shape[0]).astype(np.int32)
    ch[0] = 0
    return ch


def oracle_estep(pb, fwbw_for=None, counters=False, all_dous=None, selfing=False):
    """Runs the oracle E-step.  Returns dict(haplobase, haplocount, infprobs, homozyg, negshift, loglik, founder[, fwbw...]).
    all_dous: the analysed individuals of the whole run when pb.dous is a sample of them (children counts)."""
    K, M, C, nd = pb.K, pb.M, pb.C, len(pb.dous)
    out = dict(
        haplobase=np.zeros((K, M)), haplocount=np.zeros((K, M)), infprobs=np.zeros((K, M, 2, 2)),
        homozyg=np.zeros((K, M, 2)), negshift=np.zeros((K, M)), loglik=np.zeros((nd, C, 2)), founder=pb.founder.copy())
    children = doit_children(pb.par, pb.dous if all_dous is None else all_dous)
    cnt = np.zeros(2, np.int64) if counters else None
    fj, fc = (-1, -1) if fwbw_for is None else fwbw_for
    fw = ff = fm = None
    if fwbw_for is not None:
        mc = int(pb.chromstarts[fc + 1] - pb.chromstarts[fc])
        fw = np.zeros((8, mc, 3, 256 if selfing else 64))
        ff = np.zeros((8, mc, 3))
        fm = np.zeros(8, np.int32)
    rc = lib(selfing).cnf2_oracle_estep(
        ctypes.c_int(K), ctypes.c_int(M), ctypes.c_int(C), ctypes.c_int(nd), _p(pb.par), _p(pb.empty),
        _p(out["founder"]), _p(pb.gen), _p(pb.descendants), _p(pb.dous), _p(pb.markerpos), _p(pb.chromstarts),
        _p(pb.genrec), _p(pb.alleles), _p(pb.sure), _p(pb.haploweight), _p(out["haplobase"]), _p(out["haplocount"]),
        _p(out["infprobs"]), _p(out["homozyg"]), _p(out["loglik"]), _p(cnt), ctypes.c_int(fj), ctypes.c_int(fc),
        _p(fw), _p(ff), _p(fm), _p(out["negshift"]), _p(children))
    assert rc == 0
    if counters:
        out["counters"] = cnt
    if fwbw_for is not None:
        out.update(fwbw=fw, fwbwfactors=ff, fwbwmask=fm)
    return out


def oracle_mstep(chromstarts, active, empty, children, descendants, alleles, sure, haploweight, relhaplo, hasprior,
                 prioralleles, priorsure, haplobase, haplocount, infprobs, homozyg, scalefactor, entropyfactor):
    """One pass of the reference's M-step block (cnF2freq.cpp:6356-6367) on copies of the arrays.  Returns
    dict(alleles, sure, haploweight, haplobase, haplocount, hitnnn)."""
    K, M = haploweight.shape
    cs = np.ascontiguousarray(chromstarts, np.uint32)
    c = lambda a, dt: np.array(a, dt, order="C", copy=True)
    al, su, hw = c(alleles, np.uint8), c(sure, np.float64), c(haploweight, np.float64)
    hb, hc, ip, hz = c(haplobase, np.float64), c(haplocount, np.float64), c(infprobs, np.float64), c(homozyg, np.float64)
    f = lib().cnf2_oracle_mstep
    f.restype = ctypes.c_longlong
    hit = f(ctypes.c_int(K), ctypes.c_int(M), ctypes.c_int(len(cs) - 1), _p(cs), _p(c(active, np.uint8)), _p(c(empty, np.uint8)),
            _p(c(children, np.int32)), _p(c(descendants, np.int32)), _p(al), _p(su), _p(hw), _p(c(relhaplo, np.float64)),
            _p(c(hasprior, np.uint8)), _p(c(prioralleles, np.uint8)), _p(c(priorsure, np.float64)), _p(hb), _p(hc), _p(ip), _p(hz),
            ctypes.c_double(scalefactor), ctypes.c_double(entropyfactor))
    return dict(alleles=al, sure=su, haploweight=hw, haplobase=hb, haplocount=hc, hitnnn=int(hit))


def oracle_scalefactor(scalefactor, hitnnn, n_dous, oldhitnnn, oldhitnnn2):
    """The step-size control behind the pass (cnF2freq.cpp:6373-6390).  Returns (scalefactor, oldhitnnn, oldhitnnn2)."""
    f = lib().cnf2_oracle_scalefactor
    f.restype = ctypes.c_double
    o1, o2 = ctypes.c_int32(int(oldhitnnn)), ctypes.c_int32(int(oldhitnnn2))
    sf = f(ctypes.c_double(scalefactor), ctypes.c_longlong(hitnnn), ctypes.c_int(n_dous), ctypes.byref(o1), ctypes.byref(o2))
    return float(sf), o1.value, o2.value


def oracle_preprocess(chromstarts, par, exists, empty, alleles, sure, haploweight, selfing=False):
    """postmarkerdata (cnF2freq.cpp:3191-3412) on copies of the arrays as the readers left them.  Returns dict(alleles, sure,
    haploweight, descendants, children, founder, corrections)."""
    cs = np.ascontiguousarray(chromstarts, np.uint32)
    par = np.ascontiguousarray(par, np.int32)
    K, M = par.shape[0], np.asarray(haploweight).shape[1]
    c = lambda a, dt: np.array(a, dt, order="C", copy=True)
    al, su, hw = c(alleles, np.uint8), c(sure, np.float64), c(haploweight, np.float64)
    desc, ch, fo, n = np.zeros(K, np.int32), np.zeros(K, np.int32), np.zeros(K, np.uint8), ctypes.c_int32()
    rc = lib(selfing).cnf2_oracle_preprocess(ctypes.c_int(K), ctypes.c_int(M), ctypes.c_int(len(cs) - 1), _p(cs), _p(par), _p(c(exists, np.uint8)),
                                      _p(c(empty, np.uint8)), _p(al), _p(su), _p(hw), _p(desc), _p(ch), _p(fo), ctypes.byref(n))
    assert rc == 0
    return dict(alleles=al, sure=su, haploweight=hw, descendants=desc, children=ch, founder=fo, corrections=int(n.value),
                cleared=int(lib(selfing).cnf2_oracle_preprocess_cleared()))
