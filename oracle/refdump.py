"""Parser for the tagged-binary dumps written by oracle/ref_driver.cpp (test infrastructure only)."""
import numpy as np


def load(path):
    """Return {name: ndarray}.  Record = 'name dtype ndim d0 d1 ...\\n' + raw little-endian payload."""
    out = {}
    with open(path, "rb") as f:
        data = f.read()
    pos = 0
    while pos < len(data):
        nl = data.index(b"\n", pos)
        tok = data[pos:nl].decode().split()
        name, dtype, ndim = tok[0], tok[1], int(tok[2])
        dims = [int(x) for x in tok[3:3 + ndim]]
        n = int(np.prod(dims)) if dims else 1
        dt = np.dtype("<f8") if dtype == "d" else np.dtype("<i4")
        pos = nl + 1
        arr = np.frombuffer(data, dtype=dt, count=n, offset=pos).reshape(dims).copy()
        pos += n * dt.itemsize
        out[name] = arr
    return out
