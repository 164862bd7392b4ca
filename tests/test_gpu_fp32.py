"""GPU parity tests of FP32 mode (pp_create(..., fp_mode = 32), pp_mlp32.cu): north_star's second tolerance, relative
1e-5 against the reference's double-precision results (the reference's own float build does not compile,
Common.h:53-61, so its FP64 output is the yardstick).  Through the C ABI, like every other GPU test."""
import os

import numpy as np
import pytest

import oracle
from golden_util import TrainCase, rel_err
from prophet_b200 import capi, synth

pytestmark = pytest.mark.gpu
TOL32 = 1e-5


def test_fp32_epoch_on_the_reference_golden():
    c = TrainCase("si_train")
    ctx = capi.Context(0, fp_mode=32)
    try:
        ctx.set_descriptors(len(c.elements), c.G1, c.G2, c.G3, c.G4, c.rcut)
        params = c.params_at(0)
        ctx.set_network(0, c.sizes, params[0], c.mean(c.elements[0]), c.scale(c.elements[0]), node_transfer=c.node_transfer(c.elements[0]))
        ctx.set_output_scaling(c.out_mean, c.out_scale)
        ctx.add_structures(c.natoms, np.stack(c.cells), np.concatenate(c.frac), c.elem_flat(), c.targets, c.tags)
        for it in range(c.n_iter):
            r = ctx.epoch(c.d["x_before.%d" % it])
            assert ctx.epoch_form() == 4
            assert abs(r["SSE"] - c.d["SSE.%d" % it][0]) <= TOL32 * abs(c.d["SSE.%d" % it][0])
            assert abs(r["vSSE"] - c.d["vSSE.%d" % it][0]) <= TOL32 * max(abs(c.d["vSSE.%d" % it][0]), 1e-300)
            assert r["nval"] == int(c.d["nval.%d" % it][0])
            err = rel_err(r["gradient"], c.d["gradient.%d" % it])
            assert err < TOL32, (it, err)
            assert err > 1e-12          # it really is the single-precision kernel
        # descriptors and evaluation keep the FP64 kernels in this mode
        for s in range(c.n):
            assert rel_err(ctx.get_descriptors(s), c.G[s]) < 1e-10
    finally:
        ctx.close()


@pytest.mark.parametrize("hidden", [[40, 40], [40], [64, 24], [16, 8]])
@pytest.mark.parametrize("reps", [(2, 2, 2), (1, 1, 2), (1, 2, 3), (2, 2, 4)])
def test_fp32_epoch_shapes_and_tile_fillings_against_oracle(hidden, reps):
    """One / two hidden layers, 5- and 8-tile widths; 64-, 16-, 48- and 128-atom structures (four, sixteen, five and two
    per 256-row tile, with padding rows that the tensor copy fills with the next structure's atoms), every tag."""
    cfg = synth.config_c2(7)
    cfg.lattice = synth.lattice_si(reps)
    cfg.hidden = hidden
    n = 7
    structs = [synth.make_structure(cfg, 300 + s) for s in range(n)]
    tags = np.array([0, 0, 1, 0, 2, 0, 1], dtype=np.int32)
    net = synth.random_network(cfg, 14, 6)
    mean, scale = 0.2 * synth.normal(14, cfg.NG, 3), 0.5 + np.abs(synth.normal(14, cfg.NG, 4))
    fracs = [capi.cart_to_frac(c, x) for c, x, _, _, _ in structs]
    natoms = [len(z) for _, _, z, _, _ in structs]
    targets = np.array([t for _, _, _, t, _ in structs])
    ctx = capi.Context(0, fp_mode=32)
    try:
        ctx.set_descriptors(1, cfg.G1, cfg.G2, cfg.G3, cfg.G4, cfg.rcut)
        ctx.set_network(0, net["sizes"], net["params"], mean, scale)
        ctx.add_structures(natoms, np.stack([s[0] for s in structs]), np.concatenate(fracs), np.zeros(sum(natoms), dtype=np.int32), targets, tags)
        r = ctx.epoch(net["params"])
        # the widest shape (64 + 24) does not fit the FP32 kernel's shared-memory plan with 256-row tiles: the mode then
        # keeps the FP64 kernels, which is inside the tolerance a fortiori
        assert ctx.epoch_form() == (4 if max(hidden) <= 40 else ctx.epoch_form())
        fp32_ran = ctx.epoch_form() == 4
        G = np.concatenate([ctx.get_descriptors(s).T for s in range(n)])
        off = np.concatenate([[0], np.cumsum(natoms)]).astype(np.int32)
        ref = oracle.epoch(off, np.zeros(sum(natoms), dtype=np.int32), G, targets, tags, [oracle.Net(net["sizes"], net["params"], mean, scale)])
        assert abs(r["SSE"] - ref["SSE"]) <= TOL32 * abs(ref["SSE"])
        assert abs(r["vSSE"] - ref["vSSE"]) <= TOL32 * abs(ref["vSSE"])
        assert r["nval"] == ref["nval"] == 2
        assert rel_err(r["gradient"], ref["gradient"]) < TOL32, rel_err(r["gradient"], ref["gradient"])
        # a changed input scaling must reach the FP32 copy of the descriptors
        mean2 = mean + 0.05
        ctx.set_input_scaling(0, mean2, scale)
        r2 = ctx.epoch(net["params"])
        ref2 = oracle.epoch(off, np.zeros(sum(natoms), dtype=np.int32), G, targets, tags, [oracle.Net(net["sizes"], net["params"], mean2, scale)])
        assert rel_err(r2["gradient"], ref2["gradient"]) < TOL32
        # the same context can be told to run the FP64 kernels (diagnostic switch): the two modes agree to the FP32 tolerance
        os.environ["PP_FP32_OFF"] = "1"
        try:
            ctx.set_input_scaling(0, mean, scale)
            r64 = ctx.epoch(net["params"])
            assert ctx.epoch_form() in (0, 1, 2, 3, 5)
            assert fp32_ran or np.array_equal(r64["gradient"], r["gradient"])
        finally:
            os.environ.pop("PP_FP32_OFF", None)
        assert rel_err(r64["gradient"], ref["gradient"]) < 1e-10
    finally:
        ctx.close()


def test_fp32_mode_keeps_fp64_for_systems_it_does_not_cover():
    """Several elements: FP32 mode falls back to the FP64 two-phase kernels (documented in include/prophet_b200.h)."""
    c = TrainCase("tio2_train")
    ctx = capi.Context(0, fp_mode=32)
    try:
        ctx.set_descriptors(len(c.elements), c.G1, c.G2, c.G3, c.G4, c.rcut)
        params = c.params_at(0)
        for e, z in enumerate(c.elements):
            ctx.set_network(e, c.sizes, params[e], c.mean(z), c.scale(z), node_transfer=c.node_transfer(z))
        ctx.set_output_scaling(c.out_mean, c.out_scale)
        ctx.add_structures(c.natoms, np.stack(c.cells), np.concatenate(c.frac), c.elem_flat(), c.targets, c.tags)
        r = ctx.epoch(c.d["x_before.0"])
        assert ctx.epoch_form() in (0, 3, 5)
        assert rel_err(r["gradient"], c.d["gradient.0"]) < 1e-10
    finally:
        ctx.close()
