"""ctypes wrapper around oracle/liboracle.so (bp_oracle.c) -- TEST INFRASTRUCTURE ONLY.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may import this
package.  The product (prophet_b200/) never does: it fails loudly when its CUDA library is missing.
"""
import ctypes
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB = None

TRANSFER_IDS = {"linear": 0, "tanh": 1, "tanh_spline": 2, "tanh_crazy": 3, "cubic": 4, "pulse": 5, "sin": 6,
                "periodic_tanh": 7, "relu": 8}

_d = ctypes.POINTER(ctypes.c_double)
_i = ctypes.POINTER(ctypes.c_int)
_ll = ctypes.POINTER(ctypes.c_longlong)


def _p(a, t):
    return a.ctypes.data_as(t) if a is not None else None


def build():
    subprocess.check_call(["make", "-s", "-C", _HERE, "liboracle.so"])


def lib():
    global _LIB
    if _LIB is None:
        path = os.path.join(_HERE, "liboracle.so")
        if not os.path.exists(path):
            build()
        _LIB = ctypes.CDLL(path)
        _LIB.orc_net_atom.restype = ctypes.c_double
        _LIB.orc_transfer.restype = ctypes.c_double
        _LIB.orc_transfer.argtypes = [ctypes.c_int, ctypes.c_double, _d]
    return _LIB


def _c(a, dtype):
    return np.ascontiguousarray(a, dtype=dtype)


def _rows(rows, width):
    a = _c(np.asarray(rows, dtype=np.float64).reshape(-1, width), np.float64)
    return a, a.shape[0]


def cart_to_frac(cell, cart):
    cell = _c(cell, np.float64); cart = _c(cart, np.float64)
    out = np.empty_like(cart)
    lib().orc_cart_to_frac(_p(cell, _d), _p(cart, _d), ctypes.c_int(len(cart)), _p(out, _d))
    return out


def neighbors_periodic(cell, frac, rcut):
    cell = _c(cell, np.float64); frac = _c(frac, np.float64)
    n = len(frac)
    L = lib()
    nT = L.orc_num_translations(_p(cell, _d), ctypes.c_double(rcut))
    trans = np.zeros((nT, 3), dtype=np.int32)
    cap_j, cap_t = n * n, n * n * nT
    neigh_offset = np.zeros(n + 1, dtype=np.int32)
    neigh_j = np.zeros(cap_j, dtype=np.int32)
    image_offset = np.zeros(cap_j + 1, dtype=np.int32)
    image_t = np.zeros(cap_t, dtype=np.int32)
    rc = L.orc_neighbors_periodic(_p(cell, _d), _p(frac, _d), ctypes.c_int(n), ctypes.c_double(rcut), _p(trans, _i),
                                  _p(neigh_offset, _i), _p(neigh_j, _i), ctypes.c_int(cap_j), _p(image_offset, _i),
                                  _p(image_t, _i), ctypes.c_int(cap_t))
    assert rc == 0
    nj = neigh_offset[-1]
    return dict(translations=trans, neigh_offset=neigh_offset, neigh_j=neigh_j[:nj].copy(),
                image_offset=image_offset[:nj + 1].copy(), image_t=image_t[:image_offset[nj]].copy())


def num_descriptors(n_elem, G1, G2, G3, G4):
    return lib().orc_num_descriptors(n_elem, len(G1), len(G2), len(G3), len(G4))


def descriptors_periodic(cell, frac, elem, n_elem, G1, G2, G3, G4, rcut):
    """Returns G[NG][n] (the reference's Structure::G layout)."""
    cell = _c(cell, np.float64); frac = _c(frac, np.float64); elem = _c(elem, np.int32)
    g1, n1 = _rows(G1, 2); g2, n2 = _rows(G2, 4); g3, n3 = _rows(G3, 5); g4, n4 = _rows(G4, 5)
    NG = lib().orc_num_descriptors(n_elem, n1, n2, n3, n4)
    out = np.zeros((NG, len(frac)), dtype=np.float64)
    rc = lib().orc_descriptors_periodic(_p(cell, _d), _p(frac, _d), _p(elem, _i), ctypes.c_int(len(frac)), ctypes.c_int(n_elem),
                                        n1, _p(g1, _d), n2, _p(g2, _d), n3, _p(g3, _d), n4, _p(g4, _d),
                                        ctypes.c_double(rcut), _p(out, _d))
    assert rc == 0
    return out


def descriptors_md(x, elem, nlocal, offsets, neigh, n_elem, G1, G2, G3, G4, rcut):
    """Returns G[nlocal][NG]."""
    x = _c(x, np.float64); elem = _c(elem, np.int32); offsets = _c(offsets, np.int64); neigh = _c(neigh, np.int32)
    g1, n1 = _rows(G1, 2); g2, n2 = _rows(G2, 4); g3, n3 = _rows(G3, 5); g4, n4 = _rows(G4, 5)
    NG = lib().orc_num_descriptors(n_elem, n1, n2, n3, n4)
    out = np.zeros((nlocal, NG), dtype=np.float64)
    lib().orc_descriptors_md(_p(x, _d), _p(elem, _i), ctypes.c_int(nlocal), _p(offsets, _ll), _p(neigh, _i), ctypes.c_int(n_elem),
                             n1, _p(g1, _d), n2, _p(g2, _d), n3, _p(g3, _d), n4, _p(g4, _d), ctypes.c_double(rcut), _p(out, _d))
    return out


def forces_md(x, elem, nlocal, offsets, neigh, n_elem, G1, G2, G3, G4, rcut, dE_dG, f=None):
    x = _c(x, np.float64); elem = _c(elem, np.int32); offsets = _c(offsets, np.int64); neigh = _c(neigh, np.int32)
    dE_dG = _c(dE_dG, np.float64)
    g1, n1 = _rows(G1, 2); g2, n2 = _rows(G2, 4); g3, n3 = _rows(G3, 5); g4, n4 = _rows(G4, 5)
    if f is None:
        f = np.zeros_like(x)
    lib().orc_forces_md(_p(x, _d), _p(elem, _i), ctypes.c_int(nlocal), _p(offsets, _ll), _p(neigh, _i), ctypes.c_int(n_elem),
                        n1, _p(g1, _d), n2, _p(g2, _d), n3, _p(g3, _d), n4, _p(g4, _d), ctypes.c_double(rcut),
                        _p(dE_dG, _d), _p(f, _d))
    return f


class Net:
    """One element's network: sizes=[NG, hidden..., 1], per-node transfer ids, flat params, input mean/scale."""

    def __init__(self, sizes, params, mean, scale, transfer="tanh", node_transfer=None):
        self.sizes = _c(sizes, np.int32)
        self.params = _c(params, np.float64)
        self.mean = _c(mean, np.float64)
        self.scale = _c(scale, np.float64)
        n_nodes = int(self.sizes[1:].sum())
        if node_transfer is None:
            node_transfer = [transfer] * (n_nodes - 1) + ["linear"]
        self.tf = _c([TRANSFER_IDS[str(t)] for t in node_transfer], np.int32)
        assert len(self.tf) == n_nodes
        assert len(self.params) == sum((a + 1) * b for a, b in zip(self.sizes[:-1], self.sizes[1:]))

    @property
    def n_layers(self):
        return len(self.sizes) - 1


def net_atom(net, G, out_scale=1.0, want_dEdG=False, want_dYdP=False):
    G = _c(G, np.float64)
    dEdG = np.zeros(len(G)) if want_dEdG else None
    dYdP = np.zeros(len(net.params)) if want_dYdP else None
    e = lib().orc_net_atom(_p(net.sizes, _i), ctypes.c_int(net.n_layers), _p(net.tf, _i), _p(net.params, _d), _p(net.mean, _d),
                           _p(net.scale, _d), ctypes.c_double(out_scale), _p(G, _d), _p(dEdG, _d), _p(dYdP, _d))
    return e, dEdG, dYdP


def epoch(atom_offset, elem, G, target, tag, nets, out_mean=0.0, out_scale=1.0):
    """nets: list of Net, one per element index.  G is atom-major [total_atoms][NG]."""
    atom_offset = _c(atom_offset, np.int32); elem = _c(elem, np.int32); G = _c(G, np.float64)
    target = _c(target, np.float64); tag = _c(tag, np.int32)
    n_struct = len(target)
    NG = G.shape[1]
    n_elem = len(nets)
    max_layers = max(n.n_layers for n in nets)
    net_sizes = np.zeros((n_elem, max_layers + 1), dtype=np.int32)
    for e, n in enumerate(nets):
        net_sizes[e, :len(n.sizes)] = n.sizes
    n_layers = _c([n.n_layers for n in nets], np.int32)
    tf_all = np.concatenate([n.tf for n in nets]).astype(np.int32)
    tf_off = np.zeros(n_elem + 1, dtype=np.int32); tf_off[1:] = np.cumsum([len(n.tf) for n in nets])
    params_all = np.concatenate([n.params for n in nets])
    param_off = np.zeros(n_elem + 1, dtype=np.int32); param_off[1:] = np.cumsum([len(n.params) for n in nets])
    mean_all = _c(np.stack([n.mean for n in nets]), np.float64)
    scale_all = _c(np.stack([n.scale for n in nets]), np.float64)
    E_struct = np.zeros(n_struct); E_atom = np.zeros(len(elem))
    SSE = ctypes.c_double(); vSSE = ctypes.c_double(); nval = ctypes.c_int()
    grad = np.zeros(len(params_all))
    lib().orc_epoch(ctypes.c_int(n_struct), _p(atom_offset, _i), _p(elem, _i), _p(G, _d), ctypes.c_int(NG), _p(target, _d), _p(tag, _i),
                    ctypes.c_int(n_elem), _p(n_layers, _i), _p(net_sizes, _i), ctypes.c_int(max_layers), _p(tf_all, _i), _p(tf_off, _i),
                    _p(params_all, _d), _p(param_off, _i), _p(mean_all, _d), _p(scale_all, _d), ctypes.c_double(out_mean),
                    ctypes.c_double(out_scale), _p(E_struct, _d), ctypes.byref(SSE), ctypes.byref(vSSE), ctypes.byref(nval),
                    _p(grad, _d), _p(E_atom, _d))
    return dict(E_struct=E_struct, SSE=SSE.value, vSSE=vSSE.value, nval=nval.value, gradient=grad, E_atom=E_atom)


def precondition(atom_offset, elem, G, tag, n_elem):
    atom_offset = _c(atom_offset, np.int32); elem = _c(elem, np.int32); G = _c(G, np.float64); tag = _c(tag, np.int32)
    NG = G.shape[1]
    mean = np.zeros((n_elem, NG)); scale = np.zeros((n_elem, NG))
    lib().orc_precondition(ctypes.c_int(len(tag)), _p(atom_offset, _i), _p(elem, _i), _p(G, _d), ctypes.c_int(NG), _p(tag, _i),
                           ctypes.c_int(n_elem), _p(mean, _d), _p(scale, _d))
    return mean, scale


class Rprop:
    """State holder around orc_rprop_step (Optimizer::rprop + the history update_network keeps)."""

    def __init__(self, P):
        self.P = P
        self.weights = np.full(P, 1e-4)
        self.step = np.zeros(P)
        self.dir = np.zeros(P)
        self.dir_ready = ctypes.c_int(0)
        self.hist = []

    def update(self, x, g):
        x = _c(x, np.float64).copy(); g = _c(g, np.float64)
        self.hist.append(g.copy())
        g1 = self.hist[-2] if len(self.hist) >= 2 else g
        g2 = self.hist[-3] if len(self.hist) >= 3 else g
        lib().orc_rprop_step(ctypes.c_int(self.P), _p(x, _d), _p(g, _d), _p(g1, _d), _p(g2, _d), ctypes.c_int(len(self.hist)),
                             _p(self.weights, _d), _p(self.step, _d), _p(self.dir, _d), ctypes.byref(self.dir_ready))
        if len(self.hist) > 20:
            self.hist.pop(0)
        return x


def transfer(name, x):
    d = ctypes.c_double()
    v = lib().orc_transfer(TRANSFER_IDS[name], ctypes.c_double(x), ctypes.byref(d))
    return v, d.value
