"""GPU parity tests of the FP32 mode's tensor-core kernel (pp_mlp32u.cu: tcgen05.mma kind::tf32 for the forward and delta
products, TMEM accumulators, 128-atom tiles) beyond the shapes tests/test_gpu_fp32.py walks: both accumulator widths
(48 and 64 columns), one and two hidden layers, descriptor counts that are not multiples of four, tiles made of many
small structures, structures too large for its tiles (which keep the mma.sync kernel of pp_mlp32.cu), and the two FP32
kernels against each other.  Tolerance: north_star's relative 1e-5 against the FP64 oracle (Common.h:53-61 is the
reference's REAL switch; its own float build does not compile)."""
import os

import numpy as np
import pytest

import oracle
from golden_util import rel_err
from prophet_b200 import capi, synth

pytestmark = pytest.mark.gpu
TOL32 = 1e-5


def _system(reps, hidden, n, g2_rows=None, seed0=400):
    cfg = synth.config_c2(n)
    cfg.lattice = synth.lattice_si(reps)
    cfg.hidden = hidden
    if g2_rows is not None:
        cfg.G2 = cfg.G2[:g2_rows]
    structs = [synth.make_structure(cfg, seed0 + s) for s in range(n)]
    net = synth.random_network(cfg, 14, 9)
    mean, scale = 0.2 * synth.normal(14, cfg.NG, 3), 0.5 + np.abs(synth.normal(14, cfg.NG, 4))
    return cfg, structs, net, mean, scale


def _epoch(cfg, structs, net, mean, scale, tags, fp_mode=32):
    fracs = [capi.cart_to_frac(c, x) for c, x, _, _, _ in structs]
    natoms = [len(z) for _, _, z, _, _ in structs]
    targets = np.array([t for _, _, _, t, _ in structs])
    ctx = capi.Context(0, fp_mode=fp_mode)
    try:
        ctx.set_descriptors(1, cfg.G1, cfg.G2, cfg.G3, cfg.G4, cfg.rcut)
        ctx.set_network(0, net["sizes"], net["params"], mean, scale)
        ctx.add_structures(natoms, np.stack([s[0] for s in structs]), np.concatenate(fracs), np.zeros(sum(natoms), dtype=np.int32), targets, tags)
        r = ctx.epoch(net["params"])
        form = ctx.epoch_form()
        r2 = ctx.epoch(net["params"])
        assert np.array_equal(r["gradient"], r2["gradient"]) and r["SSE"] == r2["SSE"]      # bit-reproducible
        G = np.concatenate([ctx.get_descriptors(s).T for s in range(len(structs))])
    finally:
        ctx.close()
    off = np.concatenate([[0], np.cumsum(natoms)]).astype(np.int32)
    ref = oracle.epoch(off, np.zeros(sum(natoms), dtype=np.int32), G, targets, tags, [oracle.Net(net["sizes"], net["params"], mean, scale)])
    return r, form, ref


def _check(r, ref):
    assert abs(r["SSE"] - ref["SSE"]) <= TOL32 * abs(ref["SSE"])
    assert abs(r["vSSE"] - ref["vSSE"]) <= TOL32 * max(abs(ref["vSSE"]), 1e-300)
    assert r["nval"] == ref["nval"]
    assert rel_err(r["gradient"], ref["gradient"]) < TOL32, rel_err(r["gradient"], ref["gradient"])


@pytest.mark.parametrize("hidden", [[47, 33], [48, 40], [56, 48], [63], [48], [9, 47]])
def test_accumulator_widths_and_layer_counts_against_oracle(hidden):
    """47 is the widest layer of the 48-column accumulator, 48..63 take the 64-column one; one and two hidden layers."""
    cfg, structs, net, mean, scale = _system((2, 2, 2), hidden, 5)
    tags = np.array([0, 1, 0, 0, 2], dtype=np.int32)
    r, form, ref = _epoch(cfg, structs, net, mean, scale, tags)
    # 56 + 48 does not fit the mma.sync kernel's plan, whose work table both FP32 kernels share: that shape keeps the FP64
    # kernels (inside the tolerance a fortiori); every other one runs the FP32 mode
    assert form == (3 if hidden == [56, 48] else 4)
    _check(r, ref)


@pytest.mark.parametrize("g2_rows", [9, 7, 6])
def test_descriptor_counts_that_are_not_multiples_of_four(g2_rows):
    """NG = 49, 47, 46: the FP32 rows in HBM are padded to a 16-byte multiple and the bias column moves inside a chunk."""
    cfg, structs, net, mean, scale = _system((2, 2, 2), [40, 40], 4, g2_rows=g2_rows)
    assert cfg.NG == 40 + g2_rows
    tags = np.array([0, 0, 1, 0], dtype=np.int32)
    r, form, ref = _epoch(cfg, structs, net, mean, scale, tags)
    assert form == 4
    _check(r, ref)


def test_tiles_of_many_small_structures_and_a_partial_last_tile():
    """8-atom structures: sixteen per 128-atom tile, 43 structures -> two full tiles and a third with 88 rows; every tag."""
    cfg, structs, net, mean, scale = _system((1, 1, 1), [40, 40], 43)
    tags = np.array([(0, 0, 1, 0, 2)[s % 5] for s in range(43)], dtype=np.int32)
    r, form, ref = _epoch(cfg, structs, net, mean, scale, tags)
    assert form == 4
    _check(r, ref)


def test_structures_larger_than_the_tensor_core_tiles_keep_the_other_fp32_kernel():
    """160-atom structures do not fit 128-atom tiles: the mode stays FP32 on the mma.sync kernel (256-atom tiles)."""
    cfg, structs, net, mean, scale = _system((2, 2, 5), [40, 40], 3)
    tags = np.array([0, 1, 0], dtype=np.int32)
    r, form, ref = _epoch(cfg, structs, net, mean, scale, tags)
    assert form == 4
    _check(r, ref)


def test_the_two_fp32_kernels_agree():
    """PP_FP32_NO_UMMA=1 selects the mma.sync kernel: same 3xTF32 arithmetic, different accumulation order."""
    cfg, structs, net, mean, scale = _system((2, 2, 2), [40, 40], 6)
    tags = np.array([0, 0, 1, 0, 2, 0], dtype=np.int32)
    r, form, ref = _epoch(cfg, structs, net, mean, scale, tags)
    os.environ["PP_FP32_NO_UMMA"] = "1"
    try:
        r0, form0, _ = _epoch(cfg, structs, net, mean, scale, tags)
    finally:
        os.environ.pop("PP_FP32_NO_UMMA", None)
    assert form == 4 and form0 == 4
    _check(r, ref)
    _check(r0, ref)
    assert rel_err(r["gradient"], r0["gradient"]) < 2e-6
    assert not np.array_equal(r["gradient"], r0["gradient"])      # two different kernels did run
