"""GPU parity tests of epoch form 5: the two-phase epoch of a system of several elements through the two-CTA kernels
(mlp_fused2_kernel<FORWARD> / <TRAIN>, pp_mlp.cu) with the hidden activations H_l and their exact derivatives T'_l kept
in HBM between the two launches instead of the forward pass being run a second time (Neural_network::train_MD,
NN.cpp:626-697; Potential::train, Potential.cpp:626-670).  Through the C ABI, against the reference's golden vectors,
the pinned oracle and the recomputing form, at relative 1e-10."""
import os

import numpy as np
import pytest

import oracle
from golden_util import TrainCase, rel_err
from prophet_b200 import capi, synth
from test_gpu_stored import _build, _check, _no_store, _oracle_epoch

pytestmark = pytest.mark.gpu
TOL = 1e-10


@pytest.fixture(autouse=True)
def _default_form_selection():
    os.environ.pop("PP_MLP_FORCE_STORE", None)      # test_gpu_stored's own fixture sets it for that module only
    yield


def test_reference_golden_c3_shape_runs_form_5_and_matches_the_recompute_form():
    """tio2_train: C3's exact shape (52-40-40-1 x 2 elements) on the reference's own outputs, every recorded iteration."""
    c = TrainCase("tio2_train")
    ctx = capi.Context(0)
    try:
        ctx.set_descriptors(len(c.elements), c.G1, c.G2, c.G3, c.G4, c.rcut)
        params = c.params_at(0)
        for e, z in enumerate(c.elements):
            ctx.set_network(e, c.sizes, params[e], c.mean(z), c.scale(z), node_transfer=c.node_transfer(z))
        ctx.set_output_scaling(c.out_mean, c.out_scale)
        ctx.add_structures(c.natoms, np.stack(c.cells), np.concatenate(c.frac), c.elem_flat(), c.targets, c.tags)
        for it in range(c.n_iter):
            x = c.d["x_before.%d" % it]
            r = ctx.epoch(x)
            assert ctx.epoch_form() == 5
            assert abs(r["SSE"] - c.d["SSE.%d" % it][0]) <= TOL * abs(c.d["SSE.%d" % it][0])
            assert rel_err(r["gradient"], c.d["gradient.%d" % it]) < TOL
            r0, form0 = _no_store(ctx, x)
            assert form0 == 0
            # the stored derivatives are the ones the recomputing form parks in its scratch: same arithmetic, same order
            assert np.array_equal(r["gradient"], r0["gradient"]) and r["SSE"] == r0["SSE"]
            ra = ctx.epoch(x)
            assert np.array_equal(ra["gradient"], r["gradient"]) and ra["SSE"] == r["SSE"]      # bit-reproducible
        E = ctx.evaluate()
        mask = c.tags != 2
        assert rel_err(E[mask], c.d["E_struct.%d" % (c.n_iter - 1)][mask]) < TOL
    finally:
        ctx.close()


@pytest.mark.parametrize("reps,n,hidden", [((2, 2, 3), 5, [40, 40]), ((1, 1, 1), 3, [40, 40]), ((2, 2, 4), 1, [40, 40]),
                                            ((2, 2, 2), 4, [24]), ((2, 2, 2), 3, [64, 56]), ((2, 2, 2), 3, [33, 17]), ((2, 2, 2), 2, [7, 40])])
def test_two_element_systems_tile_fillings_and_widths_against_oracle(reps, n, hidden):
    """Row counts per element that are not multiples of the 64-row tiles, down to fewer rows than one tile and fewer tiles
    than CTAs; one and two hidden layers, five and eight n-tiles, odd widths (the store then moves 8 bytes at a time)."""
    cfg = synth.config_c3(n)
    cfg.lattice = synth.lattice_tio2(reps)
    cfg.hidden = hidden
    structs = [synth.make_structure(cfg, 700 + s) for s in range(n)]
    nets = [synth.random_network(cfg, z, 29) for z in cfg.elements]
    scal = [(0.1 * synth.normal(z, cfg.NG, 3), 0.5 + np.abs(synth.normal(z, cfg.NG, 4))) for z in cfg.elements]
    tags = [0, 1, 0, 2, 0][:n]
    ctx, natoms, elem, targets, tags = _build(cfg, structs, nets, scal, tags=tags)
    try:
        x = np.concatenate([nn["params"] for nn in nets])
        r = ctx.epoch(x)
        form = ctx.epoch_form()
        assert form == 5 or (max(hidden) > 40 and form in (0, 3))      # wide layers may not fit two CTAs per SM
        ref = _oracle_epoch(ctx, cfg, natoms, elem, targets, tags, nets, scal)
        _check(r, ref)
        if form == 5:
            r0, form0 = _no_store(ctx, x)
            assert form0 == 0 and np.array_equal(r["gradient"], r0["gradient"])
    finally:
        ctx.close()


def test_saturated_nodes_keep_their_exact_derivative_through_the_store():
    """Large weights saturate most hidden nodes: T' = 4e/(1+e)^2 is 1e-30 and smaller where 1 - H^2 would be 0 or 2e-16.
    Rprop follows the SIGN of such a gradient, so the stored T' must be the exact one."""
    cfg = synth.config_c3(3)
    cfg.lattice = synth.lattice_tio2((2, 2, 2))
    structs = [synth.make_structure(cfg, 90 + s) for s in range(3)]
    nets = [synth.random_network(cfg, z, 43) for z in cfg.elements]
    for nn in nets:
        nn["params"] = nn["params"] * 40.0
    scal = [(0.1 * synth.normal(z, cfg.NG, 3), 0.5 + np.abs(synth.normal(z, cfg.NG, 4))) for z in cfg.elements]
    ctx, natoms, elem, targets, tags = _build(cfg, structs, nets, scal)
    try:
        x = np.concatenate([nn["params"] for nn in nets])
        r = ctx.epoch(x)
        assert ctx.epoch_form() == 5
        ref = _oracle_epoch(ctx, cfg, natoms, elem, targets, tags, nets, scal)
        _check(r, ref)
        g, go = r["gradient"], ref["gradient"]
        big = np.abs(go) > 1e-200
        assert np.array_equal(np.sign(g[big]), np.sign(go[big]))
    finally:
        ctx.close()


def test_tanh_spline_networks_and_an_absent_element():
    """The reference's default transfer function through the stored form, on structures that hold only one of the two
    elements (a rank of a sharded run, Setup.cpp:582): the absent element's gradient block is exactly zero."""
    cfg = synth.config_c3(3)
    cfg.lattice = synth.lattice_tio2((2, 2, 2))
    structs = []
    for s in range(3):
        cell, cart, Z, t, tag = synth.make_structure(cfg, 60 + s)
        structs.append((cell, cart, np.full_like(Z, 8), t, tag))
    nets = [synth.random_network(cfg, z, 41) for z in cfg.elements]
    scal = [(0.1 * synth.normal(z, cfg.NG, 3), 0.5 + np.abs(synth.normal(z, cfg.NG, 4))) for z in cfg.elements]
    ctx, natoms, elem, targets, tags = _build(cfg, structs, nets, scal, transfer="tanh_spline")
    try:
        x = np.concatenate([nn["params"] for nn in nets])
        r = ctx.epoch(x)
        assert ctx.epoch_form() == 5
        P0 = len(nets[0]["params"])
        assert np.all(r["gradient"][P0:] == 0.0)
        _check(r, _oracle_epoch(ctx, cfg, natoms, elem, targets, tags, nets, scal, transfer="tanh_spline"))
    finally:
        ctx.close()


@pytest.mark.parametrize("name", ["si_train", "tio2_train", "abo4_train"])
def test_normalised_descriptor_copy_changes_nothing(name):
    """The library keeps a normalised FP64 copy of the descriptors, (x - mean) * (1 / scale) -- the arithmetic the training
    kernels otherwise apply to every tile (NN.cpp:647-651) -- and refreshes it when descriptors or scalings change.
    PP_MLP_NO_PRENORM=1 switches it off: gradients, SSE and energies must be bit-identical either way, also after the
    input scaling was changed between two epochs."""
    c = TrainCase(name)
    ctx = capi.Context(0)
    try:
        ctx.set_descriptors(len(c.elements), c.G1, c.G2, c.G3, c.G4, c.rcut)
        params = c.params_at(0)
        for e, z in enumerate(c.elements):
            ctx.set_network(e, c.sizes, params[e], c.mean(z), c.scale(z), node_transfer=c.node_transfer(z))
        ctx.set_output_scaling(c.out_mean, c.out_scale)
        ctx.add_structures(c.natoms, np.stack(c.cells), np.concatenate(c.frac), c.elem_flat(), c.targets, c.tags)
        x = c.d["x_before.0"]

        def both():
            r = ctx.epoch(x)
            E = ctx.evaluate()
            os.environ["PP_MLP_NO_PRENORM"] = "1"
            try:
                r0 = ctx.epoch(x)
                E0 = ctx.evaluate()
            finally:
                os.environ.pop("PP_MLP_NO_PRENORM", None)
            assert np.array_equal(r["gradient"], r0["gradient"]) and r["SSE"] == r0["SSE"] and np.array_equal(E, E0)
            return r

        r = both()
        assert rel_err(r["gradient"], c.d["gradient.0"]) < TOL
        for e, z in enumerate(c.elements):      # a new scaling must reach the copy
            ctx.set_input_scaling(e, c.mean(z) + 0.01, c.scale(z) * 1.1)
        r2 = both()
        assert not np.array_equal(r2["gradient"], r["gradient"])
    finally:
        ctx.close()
