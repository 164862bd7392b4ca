"""Pins oracle/bp_oracle.c (the CPU restatement) against outputs of the reference's own code (tests/golden)."""
import numpy as np
import pytest

import oracle
from golden_util import MD_CASES, TRAIN_CASES, MDCase, TrainCase, rel_err

TOL = 1e-10   # north_star: relative 1e-10 in FP64 mode


@pytest.mark.parametrize("name", TRAIN_CASES)
def test_cart_to_frac_bit_exact(name):
    c = TrainCase(name)
    for s in range(c.n):
        if not c.raw_is_cart[s]:
            assert np.array_equal(c.raw[s], c.frac[s])
            continue
        got = oracle.cart_to_frac(c.cells[s], c.raw[s])
        assert np.array_equal(got, c.frac[s]), name


@pytest.mark.parametrize("name", TRAIN_CASES)
def test_neighbor_lists_bit_exact(name):
    c = TrainCase(name)
    for s in range(c.n):
        nb = oracle.neighbors_periodic(c.cells[s], c.frac[s], c.rcut)
        assert np.array_equal(nb["translations"], c.d["translations.%d" % s])
        assert np.array_equal(nb["neigh_offset"], c.d["neigh_offset.%d" % s])
        assert np.array_equal(nb["neigh_j"], c.d["neigh_j.%d" % s])
        assert np.array_equal(nb["image_offset"], c.d["image_offset.%d" % s])
        assert np.array_equal(nb["image_t"], c.d["image_t.%d" % s])
        assert np.array_equal(np.diff(nb["neigh_offset"]), c.d["numneigh.%d" % s])


@pytest.mark.parametrize("name", TRAIN_CASES)
def test_descriptors_periodic(name):
    c = TrainCase(name)
    for s in range(c.n):
        G = oracle.descriptors_periodic(c.cells[s], c.frac[s], c.elem[s], len(c.elements), c.G1, c.G2, c.G3, c.G4, c.rcut)
        assert G.shape == c.G[s].shape
        assert rel_err(G, c.G[s]) < TOL, (name, s, rel_err(G, c.G[s]))


def _nets(c, it):
    params = c.params_at(it)
    return [oracle.Net(c.sizes, params[e], c.mean(z), c.scale(z), node_transfer=c.node_transfer(z))
            for e, z in enumerate(c.elements)]


@pytest.mark.parametrize("name", TRAIN_CASES)
def test_precondition(name):
    c = TrainCase(name)
    mean, scale = oracle.precondition(c.atom_offset, c.elem_flat(), c.G_atom_major(), c.tags, len(c.elements))
    for e, z in enumerate(c.elements):
        key = "precond_mean.%d" % z if ("precond_mean.%d" % z) in c.d else "input_mean.%d" % z
        skey = key.replace("mean", "scale")
        assert rel_err(mean[e], c.d[key]) < TOL
        assert rel_err(scale[e], c.d[skey]) < TOL


@pytest.mark.parametrize("name", TRAIN_CASES)
def test_epoch_energy_sse_gradient(name):
    c = TrainCase(name)
    G = c.G_atom_major()
    for it in range(c.n_iter):
        r = oracle.epoch(c.atom_offset, c.elem_flat(), G, c.targets, c.tags, _nets(c, it), c.out_mean, c.out_scale)
        assert rel_err(r["E_struct"], c.d["E_struct.%d" % it]) < TOL
        assert abs(r["SSE"] - c.d["SSE.%d" % it][0]) <= TOL * abs(c.d["SSE.%d" % it][0])
        assert abs(r["vSSE"] - c.d["vSSE.%d" % it][0]) <= TOL * max(abs(c.d["vSSE.%d" % it][0]), 1e-300)
        assert r["nval"] == int(c.d["nval.%d" % it][0])
        assert rel_err(r["gradient"], c.d["gradient.%d" % it]) < TOL, (name, it)
        if it == 0:
            ea = np.concatenate([c.d["E_atom.%d" % s] for s in range(c.n)])
            assert rel_err(r["E_atom"], ea) < TOL


@pytest.mark.parametrize("name", TRAIN_CASES)
def test_rprop_sequence_exact(name):
    """x_{k+1} must be reproduced EXACTLY when fed the reference's own gradients (sign-based optimizer)."""
    c = TrainCase(name)
    opt = oracle.Rprop(len(c.d["x_before.0"]))
    for it in range(c.n_iter):
        x1 = opt.update(c.d["x_before.%d" % it], c.d["gradient.%d" % it])
        assert np.array_equal(x1, c.d["x_after.%d" % it]), (name, it, np.max(np.abs(x1 - c.d["x_after.%d" % it])))


@pytest.mark.parametrize("name", MD_CASES)
def test_md_descriptors_energy_dEdG_forces(name):
    c = MDCase(name)
    ne = len(c.elements)
    G = oracle.descriptors_md(c.x, c.elem, c.nlocal, c.offsets, c.neigh, ne, c.G1, c.G2, c.G3, c.G4, c.rcut)
    assert rel_err(G, c.d["G"]) < TOL
    nets = [oracle.Net(c.sizes, c.d["params.%d" % z], c.d["in_mean.%d" % z], c.d["in_scale.%d" % z],
                       node_transfer=c.node_transfer(z)) for z in c.elements]
    eatom = np.zeros(c.nlocal)
    dEdG = np.zeros((c.nlocal, c.NG))
    for i in range(c.nlocal):
        e, d, _ = oracle.net_atom(nets[c.elem[i]], G[i], c.out_scale, want_dEdG=True)
        eatom[i] = e
        dEdG[i] = d
    assert rel_err(eatom, c.d["eatom"]) < TOL
    assert rel_err(dEdG, c.d["dE_dG"]) < TOL
    assert abs(eatom.sum() + c.out_mean - c.d["etotal"][0]) < TOL * abs(c.d["etotal"][0])
    f = oracle.forces_md(c.x, c.elem, c.nlocal, c.offsets, c.neigh, ne, c.G1, c.G2, c.G3, c.G4, c.rcut, c.d["dE_dG"])
    assert rel_err(f, c.d["f"]) < TOL, rel_err(f, c.d["f"])


def test_tanh_spline_matches_reference_table_values():
    # the variants goldens exercise tanh_spline end to end; here: the regenerated table is a C2 spline through tanh
    for x in [-20.0, -17.0, -16.99, -3.3, -0.017, 0.0, 0.5, 5.99, 6.01, 16.9, 17.0, 25.0]:
        v, d = oracle.transfer("tanh_spline", x)
        assert abs(v - np.tanh(x)) < 2e-8
        assert abs(d - (1 - np.tanh(x) ** 2)) < 2e-6
