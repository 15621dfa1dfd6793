"""The C-ABI library: it loads, exports every symbol include/cnf2_engine.h declares, and refuses to work
without a GPU (no CPU fallback).  No compute calls here."""
import ctypes
import os
import re

import pytest

from cnf2freq_b200 import engine

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def declared_symbols():
    text = open(os.path.join(ROOT, "include", "cnf2_engine.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(cnf2_[a-z_0-9]+)\s*\(", text)) - {"cnf2_allreduce_fn"})


def test_header_and_binding_agree():
    assert declared_symbols() == sorted(engine.SYMBOLS)


def test_library_exports_every_declared_symbol():
    lib = engine.load_library()
    for name in declared_symbols():
        assert hasattr(lib, name), name


def test_settings_defaults_follow_settings_h():
    lib = engine.load_library()
    s = engine.Settings()
    lib.cnf2_default_settings(ctypes.byref(s))
    assert (s.numgen, s.selfing, s.relskewstates, s.haplotyping, s.doinfprobs) == (3, 0, 0, 1, 1)


def test_unsupported_settings_are_refused():
    lib = engine.load_library()
    s = engine.Settings()
    lib.cnf2_default_settings(ctypes.byref(s))
    s.relskewstates = 1  # (SELFING = true is an instantiation of its own since row g1: tests/test_selfing_gpu.py)
    h = ctypes.c_void_p()
    assert lib.cnf2_create(ctypes.byref(h), ctypes.byref(s)) == -5
    assert b"RELSKEWSTATES" in lib.cnf2_last_error(None)
    lib.cnf2_default_settings(ctypes.byref(s))
    s.numgen = 2
    assert lib.cnf2_create(ctypes.byref(h), ctypes.byref(s)) == -5


def test_no_cpu_fallback():
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present; the refusal path is exercised on the CPU box")
    with pytest.raises(engine.EngineError, match="no CPU path|no usable CUDA device"):
        engine.Engine()


def test_host_program_builds_against_the_abi_and_refuses_to_run_without_a_gpu(tmp_path):
    """host/cnf2freq_gpu (row f4) is the compiled C++ caller of include/cnf2_engine.h: it must build with g++ against the
    header alone, parse the reference's input formats, and -- like the library -- have no CPU path."""
    import subprocess
    import torch
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    subprocess.check_call(["make", "-C", os.path.join(root, "host"), "cnf2freq_gpu"], stdout=subprocess.DEVNULL)
    data = os.path.join(root, "tests", "golden", "host", "refdata")
    r = subprocess.run([os.path.join(root, "host", "cnf2freq_gpu"), "--mapfile", data + "/demoplantimpute.map", "--pedfile",
                        data + "/demoplantimpute.ped", "--genfile", data + "/demoplantimpute.gen", "--count", "1", "--output",
                        str(tmp_path / "o")], capture_output=True, text=True)
    assert "5/1 turned into 1 1 with 0.002121;0.474640" in r.stdout  # the read-count genotypes of the demo (:6599-6657)
    if torch.cuda.is_available():
        assert r.returncode == 0
    else:
        assert r.returncode != 0 and "no usable CUDA device" in r.stderr
