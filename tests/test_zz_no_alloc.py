"""GPU test of pp_column_sums_static, the entry point for callers that may not allocate device memory.

Such callers (a host program whose own memory pool already owns the GPU, or a loop that cannot afford the
device-wide synchronisation of cudaMalloc/cudaFree) get the two reductions behind input preconditioning computed
with a fixed scratch array inside the library.  Here the entry point is called through the C ABI on descriptors
held in a caller-owned device buffer, twice in a row (the scratch is shared and keeps the previous call's
contents, so the second call proves nothing stale leaks through), and its results are compared with what the
existing allocating entry point pp_precondition installs (pp_precondition is itself pinned to the reference's
Potential::Precondition output by tests/test_gpu_parity.py::test_precondition, so agreeing with it is agreeing with
the reference) and with a plain numpy evaluation of the same sums.  The file name sorts last so that a run stopping
at the first failure has already executed every other test.
"""
import ctypes

import numpy as np
import pytest

from golden_util import TrainCase, rel_err
from prophet_b200 import capi

pytestmark = pytest.mark.gpu
TOL = 1e-10


def _sums(L, dev_matrix, n_rows, n_cols, shift, squared, out):
    rc = L.pp_column_sums_static(ctypes.c_int(0), ctypes.c_void_p(dev_matrix.data_ptr()), ctypes.c_longlong(n_rows),
                                 ctypes.c_int(n_cols), ctypes.c_void_p(shift.data_ptr()) if shift is not None else None,
                                 ctypes.c_int(squared), ctypes.c_void_p(out.data_ptr()), None)
    if rc != 0:
        raise capi.PPError(L.pp_last_error().decode())


def test_static_scratch_matches_precondition_twice():
    import torch
    L = capi.lib()
    c = TrainCase("si_train")
    # reference values from the existing (allocating) entry point
    ctx = capi.Context(0)
    try:
        ctx.set_descriptors(len(c.elements), c.G1, c.G2, c.G3, c.G4, c.rcut)
        params = c.params_at(0)
        ctx.set_network(0, c.sizes, params[0], c.mean(14), c.scale(14))
        train = [s for s in range(c.n) if c.tags[s] == 0]
        ctx.add_structures(c.natoms[train], np.stack([c.cells[s] for s in train]), np.concatenate([c.frac[s] for s in train]),
                           np.concatenate([c.elem[s] for s in train]), c.targets[train], c.tags[train])
        ctx.precondition(True)
        mean_ref, scale_ref = ctx.get_input_scaling(0)
        G = np.concatenate([ctx.get_descriptors(k).T for k in range(len(train))])      # [atoms][NG]
    finally:
        ctx.close()
    n_rows, n_cols = G.shape
    dev = torch.device("cuda", 0)
    Gd = torch.from_numpy(np.ascontiguousarray(G)).to(dev)
    sums = torch.empty(n_cols, dtype=torch.float64, device=dev)
    for call in range(2):
        # poison the output so a call that silently does nothing cannot pass
        sums.fill_(float("nan"))
        _sums(L, Gd, n_rows, n_cols, None, 0, sums)
        torch.cuda.synchronize()
        mean = sums.cpu().numpy() / n_rows
        assert rel_err(mean, mean_ref) < TOL, (call, rel_err(mean, mean_ref))
        shift = torch.from_numpy(mean).to(dev)
        sums.fill_(float("nan"))
        _sums(L, Gd, n_rows, n_cols, shift, 1, sums)
        torch.cuda.synchronize()
        scale = np.maximum(np.sqrt(sums.cpu().numpy() / n_rows), 1e-5)
        assert rel_err(scale, scale_ref) < TOL, (call, rel_err(scale, scale_ref))
        # a narrower matrix in between leaves stale wider rows in the shared scratch; the next round must not see them
        narrow = Gd[:, :7].contiguous()
        s7 = torch.empty(7, dtype=torch.float64, device=dev)
        _sums(L, narrow, n_rows, 7, None, 0, s7)
        torch.cuda.synchronize()
        assert rel_err(s7.cpu().numpy(), G[:, :7].sum(axis=0)) < TOL


def test_static_scratch_rejects_oversize():
    import torch
    L = capi.lib()
    dev = torch.device("cuda", 0)
    A = torch.zeros((4, 513), dtype=torch.float64, device=dev)
    out = torch.zeros(513, dtype=torch.float64, device=dev)
    with pytest.raises(capi.PPError, match="exceed the fixed scratch capacity"):
        _sums(L, A, 4, 513, None, 0, out)
