"""CPU-side checks of the C-ABI library: it loads, exports every symbol the header declares, its host-side
arithmetic (coordinate conversion) is bit-identical to the reference, and it refuses to run without a GPU."""
import ctypes
import os
import re

import numpy as np
import pytest

from golden_util import TRAIN_CASES, TrainCase
from prophet_b200 import capi

ROOT = os.path.abspath(os.path.join(os.path.dirname(__file__), ".."))


def _ensure_built():
    if not os.path.exists(capi.LIB_PATH):
        import __graft_entry__ as g
        g.build()


def test_library_exports_every_declared_symbol():
    _ensure_built()
    L = capi.lib()
    with open(os.path.join(ROOT, "include", "prophet_b200.h")) as f:
        header = f.read()
    declared = set(re.findall(r"\b(pp_[a-z0-9_]+)\s*\(", header))
    declared -= {"pp_context", "pp_transfer", "pp_tag"}
    assert declared == set(capi.SYMBOLS), (declared ^ set(capi.SYMBOLS))
    for name in declared:
        assert hasattr(L, name), name


@pytest.mark.parametrize("name", TRAIN_CASES)
def test_cart_to_frac_bit_exact(name):
    _ensure_built()
    c = TrainCase(name)
    for s in range(c.n):
        if c.raw_is_cart[s]:
            assert np.array_equal(capi.cart_to_frac(c.cells[s], c.raw[s]), c.frac[s])


def test_no_cpu_fallback():
    """Without a CUDA device the context cannot be created: the product never computes on the CPU."""
    _ensure_built()
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    with pytest.raises(capi.PPError, match="no CUDA device|CUDA"):
        capi.Context(0)
