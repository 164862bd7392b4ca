"""GPU parity tests: the CUDA path, called through the C ABI, against the reference's golden vectors and the
pinned oracle.  Tolerances follow BASELINE.json's north_star: relative 1e-10 (FP64 mode) with respect to the
max-abs of each array; neighbour lists bit-exact."""
import numpy as np
import pytest

import oracle
from golden_util import MD_CASES, TRAIN_CASES, MDCase, TrainCase, rel_err
from prophet_b200 import capi, synth

pytestmark = pytest.mark.gpu
TOL = 1e-10


def _has_g3(c):
    return len(c.G3) > 0


def make_ctx(c, it=0, with_nets=True):
    ctx = capi.Context(0)
    ctx.set_descriptors(len(c.elements), c.G1, c.G2, c.G3, c.G4, c.rcut)
    if with_nets:
        params = c.params_at(it)
        for e, z in enumerate(c.elements):
            ctx.set_network(e, c.sizes, params[e], c.mean(z), c.scale(z), node_transfer=c.node_transfer(z))
        ctx.set_output_scaling(c.out_mean, c.out_scale)
    return ctx


def add_all(ctx, c):
    ctx.add_structures(c.natoms, np.stack(c.cells), np.concatenate(c.frac), c.elem_flat(), c.targets, c.tags)


@pytest.fixture(scope="module", params=TRAIN_CASES)
def train_case(request):
    c = TrainCase(request.param)
    ctx = make_ctx(c)
    add_all(ctx, c)
    yield c, ctx
    ctx.close()


def test_neighbor_lists_bit_exact(train_case):
    c, ctx = train_case
    for s in range(c.n):
        nb = ctx.get_neighbors(s)
        assert np.array_equal(nb["neigh_offset"], c.d["neigh_offset.%d" % s])
        assert np.array_equal(nb["neigh_j"], c.d["neigh_j.%d" % s])
        assert np.array_equal(nb["image_offset"], c.d["image_offset.%d" % s])
        assert np.array_equal(nb["image_t"], c.d["image_t.%d" % s])


def test_descriptors(train_case):
    c, ctx = train_case
    for s in range(c.n):
        G = ctx.get_descriptors(s)
        assert G.shape == c.G[s].shape
        assert rel_err(G, c.G[s]) < TOL, (c.name, s, rel_err(G, c.G[s]))


def test_precondition(train_case):
    c, ctx = train_case
    saved = [ctx.get_input_scaling(e) for e in range(len(c.elements))]
    ctx.precondition(True)
    for e, z in enumerate(c.elements):
        key = "precond_mean.%d" % z if ("precond_mean.%d" % z) in c.d else "input_mean.%d" % z
        m, s = ctx.get_input_scaling(e)
        assert rel_err(m, c.d[key]) < TOL
        assert rel_err(s, c.d[key.replace("mean", "scale")]) < TOL
    ctx.precondition(False)
    m, s = ctx.get_input_scaling(0)
    assert np.all(m == 0) and np.all(s == 1)
    for e, (m, s) in enumerate(saved):
        ctx.set_input_scaling(e, m, s)


def test_epoch_energy_sse_gradient(train_case):
    c, ctx = train_case
    for it in range(c.n_iter):
        x = c.d["x_before.%d" % it]
        r = ctx.epoch(x)
        E = ctx.evaluate()
        ref_E = c.d["E_struct.%d" % it]
        mask = c.tags != 2                       # the reference does not evaluate 'test' structures while training
        assert rel_err(E[mask], ref_E[mask]) < TOL
        assert abs(r["SSE"] - c.d["SSE.%d" % it][0]) <= TOL * abs(c.d["SSE.%d" % it][0])
        assert abs(r["vSSE"] - c.d["vSSE.%d" % it][0]) <= TOL * max(abs(c.d["vSSE.%d" % it][0]), 1e-300)
        assert r["nval"] == int(c.d["nval.%d" % it][0])
        assert rel_err(r["gradient"], c.d["gradient.%d" % it]) < TOL, (c.name, it, rel_err(r["gradient"], c.d["gradient.%d" % it]))
    ctx.set_parameters(c.d["x_before.0"])
    ctx.evaluate()
    for s in range(c.n):
        assert rel_err(ctx.atom_energies(s), c.d["E_atom.%d" % s]) < TOL


def test_both_fused_kernel_forms_match_the_reference():
    """Single-element tanh networks train through the two-CTA kernel with in-place deltas (pp_epoch_form == 2);
    PP_MLP_ONE_CTA=1 forces the one-CTA form.  Both must reproduce the reference's SSE and gradient, and agree with
    each other far inside the tolerance (they differ only in summation order and in T' = 1 - H^2 vs 4e/(1+e)^2)."""
    import os
    c = TrainCase("si_train")
    ctx = make_ctx(c)
    try:
        add_all(ctx, c)
        x = c.d["x_before.0"]
        os.environ.pop("PP_MLP_ONE_CTA", None)
        r2 = ctx.epoch(x)
        assert ctx.epoch_form() == 2
        os.environ["PP_MLP_ONE_CTA"] = "1"
        try:
            r1 = ctx.epoch(x)
            assert ctx.epoch_form() == 1
        finally:
            os.environ.pop("PP_MLP_ONE_CTA", None)
        for r in (r1, r2):
            assert abs(r["SSE"] - c.d["SSE.0"][0]) <= TOL * abs(c.d["SSE.0"][0])
            assert rel_err(r["gradient"], c.d["gradient.0"]) < TOL
        assert rel_err(r1["gradient"], r2["gradient"]) < 1e-12
    finally:
        ctx.close()


def test_saturated_tanh_nodes_keep_their_exact_derivative():
    """The gradient of a dead neuron consists of tiny derivatives T' = 4e/(1+e)^2 only, and Rprop follows its SIGN, so
    the two-CTA form must not get them from 1 - H^2 (which is 0 or 2e-16 there): it parks the exact T' of every
    activation in a per-CTA scratch that stays in L2.  With neurons driven to |x| > 20 their gradient rows must agree
    with the one-CTA form's element by element, not merely relative to the largest gradient."""
    import os
    c = TrainCase("si_train")
    x = c.d["x_before.0"].copy()
    NG, H1 = c.sizes[0], c.sizes[1]
    x[3 * (NG + 1) + NG] = 40.0                      # bias of hidden neuron 3 of layer 0: tanh saturates for every atom
    off2 = H1 * (NG + 1)
    x[off2 + 5 * (H1 + 1) + H1] = -35.0              # and neuron 5 of the second hidden layer
    ctx = make_ctx(c)
    try:
        add_all(ctx, c)
        os.environ.pop("PP_MLP_ONE_CTA", None)
        r2 = ctx.epoch(x)
        assert ctx.epoch_form() == 2
        os.environ["PP_MLP_ONE_CTA"] = "1"
        try:
            r1 = ctx.epoch(x)
            assert ctx.epoch_form() == 1
        finally:
            os.environ.pop("PP_MLP_ONE_CTA", None)
        assert rel_err(r2["gradient"], r1["gradient"]) < 1e-12 and abs(r2["SSE"] - r1["SSE"]) <= 1e-12 * abs(r1["SSE"])
        for row1, row2 in ((r1["gradient"][3 * (NG + 1):4 * (NG + 1)], r2["gradient"][3 * (NG + 1):4 * (NG + 1)]),
                           (r1["gradient"][off2 + 5 * (H1 + 1):off2 + 6 * (H1 + 1)], r2["gradient"][off2 + 5 * (H1 + 1):off2 + 6 * (H1 + 1)])):
            assert np.all(row1 != 0.0) and np.abs(row1).max() < 1e-12 * np.abs(r1["gradient"]).max()
            assert np.allclose(row2, row1, rtol=1e-9, atol=0.0)
    finally:
        ctx.close()


def test_epoch_is_bit_reproducible(train_case):
    c, ctx = train_case
    x = c.d["x_before.0"]
    a = ctx.epoch(x)
    b = ctx.epoch(x)
    assert np.array_equal(a["gradient"], b["gradient"]) and a["SSE"] == b["SSE"]


@pytest.mark.parametrize("name", MD_CASES)
def test_md_energy_forces(name):
    c = MDCase(name)
    ctx = capi.Context(0)
    try:
        ctx.set_descriptors(len(c.elements), c.G1, c.G2, c.G3, c.G4, c.rcut)
        for e, z in enumerate(c.elements):
            ctx.set_network(e, c.sizes, c.d["params.%d" % z], c.d["in_mean.%d" % z], c.d["in_scale.%d" % z],
                            node_transfer=c.node_transfer(z))
        ctx.set_output_scaling(c.out_mean, c.out_scale)
        ctx.md_set_types([0] + list(range(len(c.elements))))
        r = ctx.md_compute(c.x, c.types, c.nlocal, c.offsets, c.neigh, want_G=True, want_dEdG=True)
        assert rel_err(r["G"], c.d["G"]) < TOL
        assert rel_err(r["eatom"], c.d["eatom"]) < TOL
        assert rel_err(r["dEdG"], c.d["dE_dG"]) < TOL
        assert abs(r["etotal"] - c.d["etotal"][0]) < TOL * abs(c.d["etotal"][0])
        assert rel_err(r["f"], c.d["f"]) < TOL, rel_err(r["f"], c.d["f"])
        # forces are accumulated INTO the caller's array (pair_nn.cpp:182)
        f0 = np.ones_like(c.x)
        r2 = ctx.md_compute(c.x, c.types, c.nlocal, c.offsets, c.neigh, f=f0.copy())
        assert rel_err(r2["f"] - 1.0, c.d["f"]) < 1e-9
    finally:
        ctx.close()


# ---- fresh seeded inputs against the oracle (sizes the oracle finishes in seconds) ---------------------------
@pytest.mark.parametrize("cfg_name,seed_base", [("C2", 100), ("C3", 200), ("C4", 300)])
def test_descriptors_vs_oracle_fresh_structures(cfg_name, seed_base):
    cfg = synth.CONFIGS[cfg_name](4)
    n = 2 if cfg_name != "C4" else 1
    zmap = {z: i for i, z in enumerate(cfg.elements)}
    ctx = capi.Context(0)
    try:
        ctx.set_descriptors(len(cfg.elements), cfg.G1, cfg.G2, cfg.G3, cfg.G4, cfg.rcut)
        structs = [synth.make_structure(cfg, seed_base + s) for s in range(n)]
        fracs, elems = [], []
        for cell, cart, Z, t, tag in structs:
            fracs.append(capi.cart_to_frac(cell, cart))
            elems.append(np.array([zmap[int(z)] for z in Z], dtype=np.int32))
        ctx.add_structures([len(e) for e in elems], np.stack([s[0] for s in structs]), np.concatenate(fracs),
                           np.concatenate(elems), [s[3] for s in structs], [0] * n)
        for s in range(n):
            G = ctx.get_descriptors(s)
            Gref = oracle.descriptors_periodic(structs[s][0], fracs[s], elems[s], len(cfg.elements), cfg.G1, cfg.G2, cfg.G3,
                                               cfg.G4, cfg.rcut)
            assert rel_err(G, Gref) < TOL, (cfg_name, s, rel_err(G, Gref))
            nb = ctx.get_neighbors(s)
            ref = oracle.neighbors_periodic(structs[s][0], fracs[s], cfg.rcut)
            for k in ("neigh_offset", "neigh_j", "image_offset", "image_t"):
                assert np.array_equal(nb[k], ref[k])
    finally:
        ctx.close()


def test_md_force_properties_full_c2_rows():
    """Size-independent properties on a larger MD system (512 local atoms): total force vanishes (Newton's third
    law over locals+ghosts), forces equal -dE/dx by central differences, and equal the oracle's on a sample."""
    cfg = synth.config_c2(1)
    cfg.lattice = synth.lattice_si((4, 4, 4))
    cell, cart, Z, _, _ = synth.make_structure(cfg, 77)
    x, Zall, nlocal = synth.expand_ghosts(cell, cart, Z, cfg.rcut + 1.0)
    numneigh, offsets, flat = synth.brute_full_neighbor_list(x, nlocal, cfg.rcut + 1.0)
    types = np.ones(len(x), dtype=np.int32)
    net = synth.random_network(cfg, 14, 5)
    mean = 0.1 * synth.normal(9, cfg.NG, 3); scale = 0.5 + np.abs(synth.normal(9, cfg.NG, 4))
    ctx = capi.Context(0)
    try:
        ctx.set_descriptors(1, cfg.G1, cfg.G2, cfg.G3, cfg.G4, cfg.rcut)
        ctx.set_network(0, net["sizes"], net["params"], mean, scale)
        ctx.md_set_types([0, 0])
        r = ctx.md_compute(x, types, nlocal, offsets, flat, want_dEdG=True)
        f = r["f"]
        assert np.max(np.abs(f.sum(axis=0))) < 1e-9 * np.max(np.abs(f))
        # oracle forces from the GPU's dE/dG on the same list
        fo = oracle.forces_md(x, np.zeros(len(x), dtype=np.int32), nlocal, offsets, flat, 1, cfg.G1, cfg.G2, cfg.G3, cfg.G4,
                              cfg.rcut, r["dEdG"])
        assert rel_err(f, fo) < TOL
    finally:
        ctx.close()


def _general_rows_config():
    """Rows that do NOT qualify for the power-basis fast path: odd and non-integer xi, three lambdas, two Rc / fc types."""
    cfg = synth.config_c3(2)
    cfg.lattice = synth.lattice_tio2((2, 2, 3))
    cfg.G2 = [[6.0, 0.0, 0.3, 0.0], [5.0, 1.0, 0.8, 1.5]]
    cfg.G4 = [[6.0, 0.0, 0.02, 3.0, 1.0], [6.0, 0.0, 0.02, 2.5, -1.0], [6.0, 0.0, 0.1, 1.0, 0.5], [6.0, 0.0, 0.1, 6.0, 1.0],
              [5.0, 1.0, 0.1, 2.0, -1.0], [5.0, 1.0, 0.3, 1.7, 1.0], [5.0, 1.0, 0.3, 3.0, 1.0]]
    cfg.G1 = [[6.0, 0.0], [4.0, 1.0]]
    return cfg


def test_general_angular_path_descriptors_and_forces_vs_oracle():
    cfg = _general_rows_config()
    zmap = {z: i for i, z in enumerate(cfg.elements)}
    cell, cart, Z, _, _ = synth.make_structure(cfg, 900)
    frac = capi.cart_to_frac(cell, cart)
    elem = np.array([zmap[int(z)] for z in Z], dtype=np.int32)
    ne = len(cfg.elements)
    ctx = capi.Context(0)
    try:
        ctx.set_descriptors(ne, cfg.G1, cfg.G2, cfg.G3, cfg.G4, cfg.rcut)
        ctx.add_structures([len(Z)], cell[None], frac, elem, [0.0], [0])
        G = ctx.get_descriptors(0)
        Gref = oracle.descriptors_periodic(cell, frac, elem, ne, cfg.G1, cfg.G2, cfg.G3, cfg.G4, cfg.rcut)
        assert rel_err(G, Gref) < TOL, rel_err(G, Gref)
        # MD branch + forces on the ghost-expanded cell, weights dE/dG from random nets
        x, Zall, nlocal = synth.expand_ghosts(cell, cart, Z, cfg.rcut + 1.0)
        numneigh, offsets, flat = synth.brute_full_neighbor_list(x, nlocal, cfg.rcut + 1.0)
        elem_all = np.array([zmap[int(z)] for z in Zall], dtype=np.int32)
        for e, z in enumerate(cfg.elements):
            net = synth.random_network(cfg, z, 17)
            ctx.set_network(e, net["sizes"], net["params"], 0.1 * synth.normal(z, cfg.NG, 3), 0.5 + np.abs(synth.normal(z, cfg.NG, 4)))
        ctx.md_set_types([0] + list(range(ne)))
        r = ctx.md_compute(x, elem_all + 1, nlocal, offsets, flat, want_G=True, want_dEdG=True)
        Gmd = oracle.descriptors_md(x, elem_all, nlocal, offsets, flat, ne, cfg.G1, cfg.G2, cfg.G3, cfg.G4, cfg.rcut)
        assert rel_err(r["G"], Gmd) < TOL
        fo = oracle.forces_md(x, elem_all, nlocal, offsets, flat, ne, cfg.G1, cfg.G2, cfg.G3, cfg.G4, cfg.rcut, r["dEdG"])
        assert rel_err(r["f"], fo) < TOL, rel_err(r["f"], fo)
    finally:
        ctx.close()


@pytest.mark.parametrize("xis", [(1, 2, 4, 8, 16), (1, 2, 4, 16), (8, 16), (1, 4), (2,)])
def test_power_basis_level_sets_vs_oracle(xis):
    """Every variant of the power-basis fast path: five squaring levels with and without xi = 8 among the rows, the
    three-level tile, and lambda groups that use different levels (Structure.h:369-414 angular_term with integer xi)."""
    cfg = synth.config_c3(2)
    cfg.lattice = synth.lattice_tio2((2, 2, 3))
    cfg.G1, cfg.G3 = [], []
    cfg.G2 = [[6.0, 0.0, 0.3, 0.0]]
    cfg.G4 = []
    for n, xi in enumerate(xis):
        cfg.G4.append([6.0, 0.0, 0.02, float(xi), 1.0])
        if n % 2 == 0:
            cfg.G4.append([6.0, 0.0, 0.2, float(xi), -1.0])
    zmap = {z: i for i, z in enumerate(cfg.elements)}
    cell, cart, Z, _, _ = synth.make_structure(cfg, 901)
    frac = capi.cart_to_frac(cell, cart)
    elem = np.array([zmap[int(z)] for z in Z], dtype=np.int32)
    ne = len(cfg.elements)
    ctx = capi.Context(0)
    try:
        ctx.set_descriptors(ne, cfg.G1, cfg.G2, cfg.G3, cfg.G4, cfg.rcut)
        ctx.add_structures([len(Z)], cell[None], frac, elem, [0.0], [0])
        G = ctx.get_descriptors(0)
        Gref = oracle.descriptors_periodic(cell, frac, elem, ne, cfg.G1, cfg.G2, cfg.G3, cfg.G4, cfg.rcut)
        assert rel_err(G, Gref) < TOL, rel_err(G, Gref)
        x, Zall, nlocal = synth.expand_ghosts(cell, cart, Z, cfg.rcut + 1.0)
        numneigh, offsets, flat = synth.brute_full_neighbor_list(x, nlocal, cfg.rcut + 1.0)
        elem_all = np.array([zmap[int(z)] for z in Zall], dtype=np.int32)
        for e, z in enumerate(cfg.elements):
            net = synth.random_network(cfg, z, 17)
            ctx.set_network(e, net["sizes"], net["params"], 0.1 * synth.normal(z, cfg.NG, 3), 0.5 + np.abs(synth.normal(z, cfg.NG, 4)))
        ctx.md_set_types([0] + list(range(ne)))
        r = ctx.md_compute(x, elem_all + 1, nlocal, offsets, flat, want_G=True, want_dEdG=True)
        Gmd = oracle.descriptors_md(x, elem_all, nlocal, offsets, flat, ne, cfg.G1, cfg.G2, cfg.G3, cfg.G4, cfg.rcut)
        assert rel_err(r["G"], Gmd) < TOL
        fo = oracle.forces_md(x, elem_all, nlocal, offsets, flat, ne, cfg.G1, cfg.G2, cfg.G3, cfg.G4, cfg.rcut, r["dEdG"])
        assert rel_err(r["f"], fo) < TOL, rel_err(r["f"], fo)
    finally:
        ctx.close()


@pytest.mark.parametrize("rcut", [5.55, 5.9, 6.0])
def test_neighbour_counts_across_the_tail_pass_split(rcut):
    """The pair loops split the partner range of a short last pass over 2-4 lane groups (pp_desc.cuh).  Rcut 5.55 gives
    34 neighbours (2 left over: four groups), 5.9 cuts through the perturbed fifth shell (34..46 neighbours: every group
    count), 6.0 is C2's 46.  Descriptors (periodic and MD branch) and forces against the oracle."""
    cfg = synth.config_c2(1)
    cfg.sigma = 0.12
    cfg.rcut = rcut
    cfg.G2 = [[rcut] + list(r[1:]) for r in cfg.G2][:4]
    cfg.G4 = [[rcut] + list(r[1:]) for r in cfg.G4]
    cell, cart, Z, _, _ = synth.make_structure(cfg, 31)
    frac = capi.cart_to_frac(cell, cart)
    elem = np.zeros(len(Z), dtype=np.int32)
    ctx = capi.Context(0)
    try:
        ctx.set_descriptors(1, cfg.G1, cfg.G2, cfg.G3, cfg.G4, cfg.rcut)
        ctx.add_structures([len(Z)], cell[None], frac, elem, [0.0], [0])
        G = ctx.get_descriptors(0)
        Gref = oracle.descriptors_periodic(cell, frac, elem, 1, cfg.G1, cfg.G2, cfg.G3, cfg.G4, cfg.rcut)
        assert rel_err(G, Gref) < TOL, rel_err(G, Gref)
        x, Zall, nlocal = synth.expand_ghosts(cell, cart, Z, cfg.rcut + 1.0)
        numneigh, offsets, flat = synth.brute_full_neighbor_list(x, nlocal, cfg.rcut + 1.0)
        within = []
        for i in range(nlocal):
            js = flat[offsets[i]:offsets[i + 1]]
            d = np.linalg.norm(x[js] - x[i], axis=1)
            within.append(int(np.sum(d <= cfg.rcut)))
        if rcut == 5.9:
            assert len(set(within)) >= 5, sorted(set(within))          # the split is exercised at several widths
        net = synth.random_network(cfg, 14, 5)
        ctx.set_network(0, net["sizes"], net["params"], 0.1 * synth.normal(9, cfg.NG, 3), 0.5 + np.abs(synth.normal(9, cfg.NG, 4)))
        ctx.md_set_types([0, 0])
        types = np.ones(len(x), dtype=np.int32)
        r = ctx.md_compute(x, types, nlocal, offsets, flat, want_G=True, want_dEdG=True)
        zeros = np.zeros(len(x), dtype=np.int32)
        Gmd = oracle.descriptors_md(x, zeros, nlocal, offsets, flat, 1, cfg.G1, cfg.G2, cfg.G3, cfg.G4, cfg.rcut)
        assert rel_err(r["G"], Gmd) < TOL, rel_err(r["G"], Gmd)
        fo = oracle.forces_md(x, zeros, nlocal, offsets, flat, 1, cfg.G1, cfg.G2, cfg.G3, cfg.G4, cfg.rcut, r["dEdG"])
        assert rel_err(r["f"], fo) < TOL, rel_err(r["f"], fo)
    finally:
        ctx.close()


def test_default_auto_generated_rows_descriptors_and_forces_vs_oracle():
    """What a user who writes no G rows gets (Structure::init_G, Structure.cpp:816-850, `Nradial 8`, `Nangular 8`): G2
    rows with one eta and eight Rs, G4 rows with xi = 1, 6, 11, 16 and lambda = +-1 -- odd exponents that the power
    basis does not cover, so the general angular path runs.  Periodic descriptors and neighbour lists, then the MD
    branch and the force chain rule on the ghost-expanded cell, against the oracle."""
    from prophet_b200 import hostapi
    g2, g4 = hostapi.auto_rows(6.0, 8, 8)
    assert sorted({float(r[3]) for r in g4}) == [1.0, 6.0, 11.0, 16.0]
    cfg = synth.config_c3(2)
    cfg.lattice = synth.lattice_tio2((2, 2, 3))
    cfg.G1, cfg.G3 = [], []
    cfg.G2, cfg.G4 = [list(r) for r in g2], [list(r) for r in g4]
    zmap = {z: i for i, z in enumerate(cfg.elements)}
    cell, cart, Z, _, _ = synth.make_structure(cfg, 910)
    frac = capi.cart_to_frac(cell, cart)
    elem = np.array([zmap[int(z)] for z in Z], dtype=np.int32)
    ne = len(cfg.elements)
    ctx = capi.Context(0)
    try:
        ctx.set_descriptors(ne, cfg.G1, cfg.G2, cfg.G3, cfg.G4, cfg.rcut)
        ctx.add_structures([len(Z)], cell[None], frac, elem, [0.0], [0])
        G = ctx.get_descriptors(0)
        Gref = oracle.descriptors_periodic(cell, frac, elem, ne, cfg.G1, cfg.G2, cfg.G3, cfg.G4, cfg.rcut)
        assert rel_err(G, Gref) < TOL, rel_err(G, Gref)
        x, Zall, nlocal = synth.expand_ghosts(cell, cart, Z, cfg.rcut + 1.0)
        numneigh, offsets, flat = synth.brute_full_neighbor_list(x, nlocal, cfg.rcut + 1.0)
        elem_all = np.array([zmap[int(z)] for z in Zall], dtype=np.int32)
        for e, z in enumerate(cfg.elements):
            net = synth.random_network(cfg, z, 19)
            ctx.set_network(e, net["sizes"], net["params"], 0.1 * synth.normal(z, cfg.NG, 3), 0.5 + np.abs(synth.normal(z, cfg.NG, 4)))
        ctx.md_set_types([0] + list(range(ne)))
        r = ctx.md_compute(x, elem_all + 1, nlocal, offsets, flat, want_G=True, want_dEdG=True)
        Gmd = oracle.descriptors_md(x, elem_all, nlocal, offsets, flat, ne, cfg.G1, cfg.G2, cfg.G3, cfg.G4, cfg.rcut)
        assert rel_err(r["G"], Gmd) < TOL
        fo = oracle.forces_md(x, elem_all, nlocal, offsets, flat, ne, cfg.G1, cfg.G2, cfg.G3, cfg.G4, cfg.rcut, r["dEdG"])
        assert rel_err(r["f"], fo) < TOL, rel_err(r["f"], fo)
    finally:
        ctx.close()


def test_forty_distinct_g4_keys_the_optimize_search_on_c2_rows():
    """Potential::optimize_Gs (Potential.cpp:709-877) perturbs eta, xi of EVERY row, so C2's 40 G4 rows become 40
    distinct (Rc, fc, eta) keys and 40 distinct non-integer (xi, lambda) keys: the general angular path walks the
    whole 40 x 40 key grid in chunks.  pp_set_descriptors with new row parameters + pp_recompute_descriptors against
    the oracle, descriptors and forces."""
    cfg = synth.config_c2(1)
    rng = np.random.RandomState(3)
    g2 = [[r[0], r[1], r[2] * (1.0 + 0.1 * rng.rand()), r[3]] for r in cfg.G2]
    g4 = [[r[0], r[1], r[2] * (1.0 + 0.1 * rng.rand()), r[3] * (1.0 + 0.1 * rng.rand()), r[4]] for r in cfg.G4]
    assert len({(r[0], r[1], r[2]) for r in g4}) == 40 and len({(r[3], r[4]) for r in g4}) == 40
    cell, cart, Z, _, _ = synth.make_structure(cfg, 920)
    frac = capi.cart_to_frac(cell, cart)
    elem = np.zeros(len(Z), dtype=np.int32)
    ctx = capi.Context(0)
    try:
        ctx.set_descriptors(1, cfg.G1, cfg.G2, cfg.G3, cfg.G4, cfg.rcut)          # the power-basis rows first
        ctx.add_structures([len(Z)], cell[None], frac, elem, [0.0], [0])
        ctx.set_descriptors(1, cfg.G1, g2, cfg.G3, g4, cfg.rcut)                  # then the perturbed parameters, in place
        ctx.recompute_descriptors()
        G = ctx.get_descriptors(0)
        Gref = oracle.descriptors_periodic(cell, frac, elem, 1, cfg.G1, g2, cfg.G3, g4, cfg.rcut)
        assert rel_err(G, Gref) < TOL, rel_err(G, Gref)
        x, Zall, nlocal = synth.expand_ghosts(cell, cart, Z, cfg.rcut + 1.0)
        numneigh, offsets, flat = synth.brute_full_neighbor_list(x, nlocal, cfg.rcut + 1.0)
        net = synth.random_network(cfg, 14, 21)
        ctx.set_network(0, net["sizes"], net["params"], 0.1 * synth.normal(14, cfg.NG, 3), 0.5 + np.abs(synth.normal(14, cfg.NG, 4)))
        ctx.md_set_types([0, 0])
        r = ctx.md_compute(x, np.ones(len(x), dtype=np.int32), nlocal, offsets, flat, want_G=True, want_dEdG=True)
        fo = oracle.forces_md(x, np.zeros(len(x), dtype=np.int32), nlocal, offsets, flat, 1, cfg.G1, g2, cfg.G3, g4, cfg.rcut, r["dEdG"])
        assert rel_err(r["f"], fo) < TOL, rel_err(r["f"], fo)
    finally:
        ctx.close()


@pytest.mark.gpu
def test_gaussian_factors_chained_as_integer_powers_vs_oracle():
    """pp_set_descriptors orders the Gaussian factors of the G2 rows and of the G4 radial keys by (Rs, eta) and evaluates an
    eta that is an integer multiple of its predecessor's as a power of the predecessor's exponential (DescSpec::pair_order).
    Rows chosen to walk every branch: ratios 1 (same eta twice), 3, 4 and 64, a ratio of 65 (too large: fresh exp), a
    non-integer ratio, chains interrupted by a different Rs, rows given in descending eta, different Rc inside one chain
    (an entry beyond the smaller Rc must still feed its successor), and a G1 row in front.  Descriptors and forces against
    the oracle, which calls exp() for every row as the reference does (Structure.h:351-360)."""
    g1 = [[6.0, 0.0]]
    g2 = [[6.0, 0.0, 0.9, 1.0], [6.0, 0.0, 0.3, 0.0], [4.5, 0.0, 0.1, 0.0], [6.0, 0.0, 0.3, 1.0], [6.0, 0.0, 0.1, 1.0],
          [6.0, 0.0, 0.3, 0.0], [6.0, 0.0, 6.4, 0.0], [6.0, 0.0, 6.5 * 6.4, 2.0], [6.0, 0.0, 0.1 * 65.0, 2.0], [6.0, 0.0, 0.1, 2.0],
          [6.0, 0.0, 0.25, 2.0]]
    g4 = [[6.0, 0.0, eta, xi, lam] for eta in (0.27, 0.03, 0.09, 0.0301) for (xi, lam) in ((1.0, 1.0), (2.0, -1.0), (4.0, 1.0))]
    cfg = synth.config_c2(1)
    cell, cart, Z, _, _ = synth.make_structure(cfg, 931)
    frac = capi.cart_to_frac(cell, cart)
    elem = np.zeros(len(Z), dtype=np.int32)
    ctx = capi.Context(0)
    try:
        ctx.set_descriptors(1, g1, g2, [], g4, cfg.rcut)
        ctx.add_structures([len(Z)], cell[None], frac, elem, [0.0], [0])
        G = ctx.get_descriptors(0)
        Gref = oracle.descriptors_periodic(cell, frac, elem, 1, g1, g2, [], g4, cfg.rcut)
        assert rel_err(G, Gref) < TOL, rel_err(G, Gref)
        NG = len(g1) + len(g2) + len(g4)
        x, Zall, nlocal = synth.expand_ghosts(cell, cart, Z, cfg.rcut + 1.0)
        numneigh, offsets, flat = synth.brute_full_neighbor_list(x, nlocal, cfg.rcut + 1.0)
        rng = np.random.RandomState(5)
        sizes = [NG, 12, 1]
        params = 0.3 * rng.randn(NG * 12 + 12 + 12 + 1)
        ctx.set_network(0, sizes, params, 0.1 * synth.normal(14, NG, 3), 0.5 + np.abs(synth.normal(14, NG, 4)))
        ctx.md_set_types([0, 0])
        r = ctx.md_compute(x, np.ones(len(x), dtype=np.int32), nlocal, offsets, flat, want_G=True, want_dEdG=True)
        Gmd = oracle.descriptors_md(x, np.zeros(len(x), dtype=np.int32), nlocal, offsets, flat, 1, g1, g2, [], g4, cfg.rcut)
        assert rel_err(r["G"], Gmd) < TOL, rel_err(r["G"], Gmd)
        fo = oracle.forces_md(x, np.zeros(len(x), dtype=np.int32), nlocal, offsets, flat, 1, g1, g2, [], g4, cfg.rcut, r["dEdG"])
        assert rel_err(r["f"], fo) < TOL, rel_err(r["f"], fo)
    finally:
        ctx.close()
