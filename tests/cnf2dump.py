"""Reader for the CNF2DUMP container written by oracle/ref_driver.cpp (test infrastructure).

Records: ``ARRAY <name> <dtype> <ndim> <d0> ... <dn-1>\\n`` + raw little-endian bytes + ``\\n``.
"""
import numpy as np

_DT = {"f64": np.float64, "i32": np.int32, "u8": np.uint8}


def read_dump(path):
    out = {}
    with open(path, "rb") as f:
        data = f.read()
    pos = 0
    while pos < len(data):
        nl = data.index(b"\n", pos)
        hdr = data[pos:nl].decode().split()
        assert hdr[0] == "ARRAY", hdr
        name, dt, nd = hdr[1], _DT[hdr[2]], int(hdr[3])
        dims = [int(x) for x in hdr[4:4 + nd]]
        n = int(np.prod(dims)) if dims else 1
        nbytes = n * np.dtype(dt).itemsize
        pos = nl + 1
        out[name] = np.frombuffer(data[pos:pos + nbytes], dtype=dt).reshape(dims).copy()
        pos += nbytes + 1
    return out


def write_npz(path, **arrays):
    np.savez_compressed(path, **arrays)
