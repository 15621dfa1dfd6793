"""ctypes binding of the CPU oracle (oracle/cnf2_oracle.cpp).  TEST INFRASTRUCTURE ONLY."""
import ctypes
import os
import subprocess

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
_LIB = os.path.join(ROOT, "oracle", "_build", "libcnf2oracle.so")


def build_oracle(force=False):
    src = os.path.join(ROOT, "oracle", "cnf2_oracle.cpp")
    if force or not os.path.exists(_LIB) or os.path.getmtime(_LIB) < os.path.getmtime(src):
        os.makedirs(os.path.dirname(_LIB), exist_ok=True)
        subprocess.check_call(["g++", "-std=c++17", "-O2", "-fopenmp", "-fPIC", "-shared", "-o", _LIB, src])
    return _LIB


_lib = None


def lib():
    global _lib
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


def oracle_estep(pb, fwbw_for=None, counters=False):
    """Runs the oracle E-step.  Returns dict(haplobase, haplocount, infprobs, homozyg, loglik, founder[, fwbw...])."""
    K, M, C, nd = pb.K, pb.M, pb.C, len(pb.dous)
    out = dict(
        haplobase=np.zeros((K, M)), haplocount=np.zeros((K, M)), infprobs=np.zeros((K, M, 2, 2)),
        homozyg=np.zeros((K, M, 2)), loglik=np.zeros((nd, C, 2)), founder=pb.founder.copy())
    cnt = np.zeros(2, np.int64) if counters else None
    fj, fc = (-1, -1) if fwbw_for is None else fwbw_for
    fw = ff = fm = None
    if fwbw_for is not None:
        mc = int(pb.chromstarts[fc + 1] - pb.chromstarts[fc])
        fw = np.zeros((8, mc, 3, 64))
        ff = np.zeros((8, mc, 3))
        fm = np.zeros(8, np.int32)
    rc = lib().cnf2_oracle_estep(
        ctypes.c_int(K), ctypes.c_int(M), ctypes.c_int(C), ctypes.c_int(nd), _p(pb.par), _p(pb.empty),
        _p(out["founder"]), _p(pb.gen), _p(pb.descendants), _p(pb.dous), _p(pb.markerpos), _p(pb.chromstarts),
        _p(pb.genrec), _p(pb.alleles), _p(pb.sure), _p(pb.haploweight), _p(out["haplobase"]), _p(out["haplocount"]),
        _p(out["infprobs"]), _p(out["homozyg"]), _p(out["loglik"]), _p(cnt), ctypes.c_int(fj), ctypes.c_int(fc),
        _p(fw), _p(ff), _p(fm))
    assert rc == 0
    if counters:
        out["counters"] = cnt
    if fwbw_for is not None:
        out.update(fwbw=fw, fwbwfactors=ff, fwbwmask=fm)
    return out
