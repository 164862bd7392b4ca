"""GPU parity tests of the two-phase epoch that keeps the activations in HBM between its phases (pp_mlpw.cu,
pp_epoch_form == 3): the path systems of several elements and wide / deep networks train through
(Neural_network::train_MD, NN.cpp:626-697; Potential::train, Potential.cpp:626-670).  Everything goes through the
C ABI and is compared with the pinned oracle (or the reference's golden vectors) at relative 1e-10."""
import os

import numpy as np
import pytest

import oracle
from golden_util import TrainCase, rel_err
from prophet_b200 import capi, synth

pytestmark = pytest.mark.gpu
TOL = 1e-10


@pytest.fixture(autouse=True)
def _force_stored_form():
    """Networks the two-CTA kernels can run (C3's 52-40-40-1) default to those; these tests are about the stored form."""
    os.environ["PP_MLP_FORCE_STORE"] = "1"
    yield
    os.environ.pop("PP_MLP_FORCE_STORE", None)


def _build(cfg, structs, nets, scalings, tags=None, transfer="tanh", node_transfer=None):
    zmap = {z: i for i, z in enumerate(cfg.elements)}
    fracs = [capi.cart_to_frac(c, x) for c, x, _, _, _ in structs]
    elems = [np.array([zmap[int(z)] for z in Z], dtype=np.int32) for _, _, Z, _, _ in structs]
    natoms = [len(e) for e in elems]
    targets = np.array([s[3] for s in structs])
    tags = np.zeros(len(structs), dtype=np.int32) if tags is None else np.asarray(tags, dtype=np.int32)
    ctx = capi.Context(0)
    ctx.set_descriptors(len(cfg.elements), cfg.G1, cfg.G2, cfg.G3, cfg.G4, cfg.rcut)
    for e in range(len(cfg.elements)):
        ctx.set_network(e, nets[e]["sizes"], nets[e]["params"], scalings[e][0], scalings[e][1], transfer=transfer,
                        node_transfer=None if node_transfer is None else node_transfer[e])
    ctx.add_structures(natoms, np.stack([s[0] for s in structs]), np.concatenate(fracs), np.concatenate(elems), targets, tags)
    return ctx, natoms, np.concatenate(elems), targets, tags


def _oracle_epoch(ctx, cfg, natoms, elem, targets, tags, nets, scalings, transfer="tanh", node_transfer=None):
    G = np.concatenate([ctx.get_descriptors(s).T for s in range(len(natoms))])      # descriptors are checked elsewhere
    off = np.concatenate([[0], np.cumsum(natoms)]).astype(np.int32)
    onets = [oracle.Net(nets[e]["sizes"], nets[e]["params"], scalings[e][0], scalings[e][1], transfer=transfer,
                        node_transfer=None if node_transfer is None else node_transfer[e]) for e in range(len(nets))]
    return oracle.epoch(off, elem, G, targets, tags, onets)


def _check(r, ref):
    assert abs(r["SSE"] - ref["SSE"]) <= TOL * abs(ref["SSE"])
    assert abs(r["vSSE"] - ref["vSSE"]) <= TOL * max(abs(ref["vSSE"]), 1e-300)
    assert r["nval"] == ref["nval"]
    assert rel_err(r["gradient"], ref["gradient"]) < TOL, rel_err(r["gradient"], ref["gradient"])


def _no_store(ctx, x):
    os.environ["PP_MLP_NO_STORE"] = "1"
    try:
        r = ctx.epoch(x)
        form = ctx.epoch_form()
    finally:
        os.environ.pop("PP_MLP_NO_STORE", None)
    return r, form


@pytest.mark.parametrize("name", ["tio2_train", "abo4_train"])
def test_reference_goldens_run_the_stored_form_and_match_the_recompute_form(name):
    """C3's (52-40-40-1 x 2 elements) and C4's (120-60-60-60-1 x 3 elements) exact network shapes on the reference's
    own outputs; PP_MLP_NO_STORE=1 selects the older form that recomputes the forward pass: same arithmetic."""
    c = TrainCase(name)
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
            assert ctx.epoch_form() == 3
            assert abs(r["SSE"] - c.d["SSE.%d" % it][0]) <= TOL * abs(c.d["SSE.%d" % it][0])
            assert rel_err(r["gradient"], c.d["gradient.%d" % it]) < TOL
            r0, form0 = _no_store(ctx, x)
            assert form0 == 0
            assert rel_err(r["gradient"], r0["gradient"]) < 1e-12 and abs(r["SSE"] - r0["SSE"]) <= 1e-12 * abs(r0["SSE"])
            ra = ctx.epoch(x)
            assert np.array_equal(ra["gradient"], r["gradient"]) and ra["SSE"] == r["SSE"]      # bit-reproducible
        E = ctx.evaluate()
        mask = c.tags != 2
        assert rel_err(E[mask], c.d["E_struct.%d" % (c.n_iter - 1)][mask]) < TOL
    finally:
        ctx.close()


def test_c4_shape_fresh_structures_against_oracle():
    """BASELINE config 4 as named -- ternary, 216 atoms, 120 descriptors, three nets 120-60-60-60-1 -- on fresh seeded
    structures with a validation structure among them: SSE, vSSE, gradient and per-atom energies against the oracle."""
    cfg = synth.config_c4(4)
    structs = [synth.make_structure(cfg, 500 + s) for s in range(4)]
    nets = [synth.random_network(cfg, z, 23) for z in cfg.elements]
    scal = [(0.1 * synth.normal(z, cfg.NG, 3), 0.5 + np.abs(synth.normal(z, cfg.NG, 4))) for z in cfg.elements]
    ctx, natoms, elem, targets, tags = _build(cfg, structs, nets, scal, tags=[0, 0, 1, 0])
    try:
        assert [int(v) for v in nets[0]["sizes"]] == [120, 60, 60, 60, 1] and len(cfg.elements) == 3
        x = np.concatenate([n["params"] for n in nets])
        r = ctx.epoch(x)
        assert ctx.epoch_form() == 3
        ref = _oracle_epoch(ctx, cfg, natoms, elem, targets, tags, nets, scal)
        _check(r, ref)
        ctx.evaluate()
        off = np.concatenate([[0], np.cumsum(natoms)])
        for s in range(len(natoms)):
            assert rel_err(ctx.atom_energies(s), ref["E_atom"][off[s]:off[s + 1]]) < TOL
    finally:
        ctx.close()


@pytest.mark.parametrize("reps,n", [((2, 2, 3), 5), ((1, 1, 1), 3), ((2, 2, 4), 1)])
def test_c3_shape_tile_fillings_against_oracle(reps, n):
    """Two elements, nets 52-40-40-1 (five n-tiles: three teams of five warps).  Row counts per element that are not
    multiples of the 24-row team tiles nor of the 64-row backward tiles, down to fewer rows than one tile."""
    cfg = synth.config_c3(n)
    cfg.lattice = synth.lattice_tio2(reps)
    structs = [synth.make_structure(cfg, 700 + s) for s in range(n)]
    nets = [synth.random_network(cfg, z, 29) for z in cfg.elements]
    scal = [(0.1 * synth.normal(z, cfg.NG, 3), 0.5 + np.abs(synth.normal(z, cfg.NG, 4))) for z in cfg.elements]
    tags = [0, 1, 0, 2, 0][:n]
    ctx, natoms, elem, targets, tags = _build(cfg, structs, nets, scal, tags=tags)
    try:
        x = np.concatenate([nn["params"] for nn in nets])
        r = ctx.epoch(x)
        assert ctx.epoch_form() == 3
        _check(r, _oracle_epoch(ctx, cfg, natoms, elem, targets, tags, nets, scal))
    finally:
        ctx.close()


@pytest.mark.parametrize("hidden", [[60, 60, 60], [64, 64], [16], [48, 20, 36, 8]])
def test_single_element_wide_networks_against_oracle(hidden):
    """Single-element systems whose networks are too large for 64-atom tiles in the single-pass kernels take the stored
    form as well (one to four hidden layers, widths up to 64)."""
    cfg = synth.config_c2(3)
    cfg.hidden = hidden
    structs = [synth.make_structure(cfg, 800 + s) for s in range(3)]
    nets = [synth.random_network(cfg, 14, 31)]
    scal = [(0.2 * synth.normal(14, cfg.NG, 3), 0.5 + np.abs(synth.normal(14, cfg.NG, 4)))]
    ctx, natoms, elem, targets, tags = _build(cfg, structs, nets, scal, tags=[0, 1, 0])
    try:
        r = ctx.epoch(nets[0]["params"])
        ref = _oracle_epoch(ctx, cfg, natoms, elem, targets, tags, nets, scal)
        _check(r, ref)
        # whatever form the plans chose for the default run, the stored form must agree when it is the one running
        if ctx.epoch_form() == 3:
            r0, form0 = _no_store(ctx, nets[0]["params"])
            assert form0 in (0, 1, 2)
            assert rel_err(r["gradient"], r0["gradient"]) < 1e-12
    finally:
        ctx.close()


def test_every_transfer_function_through_the_stored_form():
    """Mixed transfer functions per node (the generic epilogue of phase 1; phase 2 only sees the stored T')."""
    names = ["tanh", "tanh_spline", "tanh_crazy", "cubic", "pulse", "sin", "periodic_tanh", "relu", "linear"]
    cfg = synth.config_c3(2)
    cfg.lattice = synth.lattice_tio2((2, 2, 2))
    cfg.hidden = [10, 8]
    structs = [synth.make_structure(cfg, 40 + s) for s in range(2)]
    nets = [synth.random_network(cfg, z, 37) for z in cfg.elements]
    scal = [(0.1 * synth.normal(z, cfg.NG, 3), 0.5 + np.abs(synth.normal(z, cfg.NG, 4))) for z in cfg.elements]
    nt = [[names[(k + 3 * e) % len(names)] for k in range(18)] + ["linear"] for e in range(2)]
    ctx, natoms, elem, targets, tags = _build(cfg, structs, nets, scal, node_transfer=nt)
    try:
        x = np.concatenate([nn["params"] for nn in nets])
        r = ctx.epoch(x)
        assert ctx.epoch_form() == 3
        _check(r, _oracle_epoch(ctx, cfg, natoms, elem, targets, tags, nets, scal, node_transfer=nt))
    finally:
        ctx.close()


@pytest.mark.parametrize("store", [True, False])
def test_element_of_the_potential_absent_from_this_ranks_structures(store):
    """A potential for two elements trained on structures that hold only one of them (what a rank of a sharded run can
    end up with, Setup.cpp:582): the absent element's gradient block is exactly zero and the other one is unaffected,
    in both two-phase forms."""
    cfg = synth.config_c3(3)
    cfg.lattice = synth.lattice_tio2((2, 2, 2))
    structs = []
    for s in range(3):
        cell, cart, Z, t, tag = synth.make_structure(cfg, 60 + s)
        structs.append((cell, cart, np.full_like(Z, 8), t, tag))            # every atom an oxygen
    nets = [synth.random_network(cfg, z, 41) for z in cfg.elements]
    scal = [(0.1 * synth.normal(z, cfg.NG, 3), 0.5 + np.abs(synth.normal(z, cfg.NG, 4))) for z in cfg.elements]
    ctx, natoms, elem, targets, tags = _build(cfg, structs, nets, scal)
    try:
        assert cfg.elements == [8, 22] and np.all(elem == 0)
        x = np.concatenate([nn["params"] for nn in nets])
        if store:
            r = ctx.epoch(x)
            assert ctx.epoch_form() == 3
        else:
            r, form = _no_store(ctx, x)
            assert form == 0
        P0 = len(nets[0]["params"])
        assert np.all(r["gradient"][P0:] == 0.0)
        _check(r, _oracle_epoch(ctx, cfg, natoms, elem, targets, tags, nets, scal))
    finally:
        ctx.close()
