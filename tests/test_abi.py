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
    s.selfing = 1
    h = ctypes.c_void_p()
    assert lib.cnf2_create(ctypes.byref(h), ctypes.byref(s)) == -5
    assert b"SELFING" in lib.cnf2_last_error(None)


def test_no_cpu_fallback():
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present; the refusal path is exercised on the CPU box")
    with pytest.raises(engine.EngineError, match="no CPU path|no usable CUDA device"):
        engine.Engine()
