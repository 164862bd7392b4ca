"""ctypes binding of the C ABI in include/prophet_b200.h (libprophet_b200.so).

This is the same binding any host language would write against the header; tests and
bench.py go through it so that what they exercise is the shipped C ABI.  There is NO fallback: if the shared
library has not been built, or the machine has no CUDA device, the calls raise.
"""
from __future__ import annotations

import ctypes
import os
from typing import Optional, Sequence

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("PP_LIB_PATH") or os.path.join(_HERE, "libprophet_b200.so")      # PP_LIB_PATH: tuning builds (tools/)
_LIB = None

TRANSFER_IDS = {"linear": 0, "tanh": 1, "tanh_spline": 2, "tanh_crazy": 3, "cubic": 4, "pulse": 5, "sin": 6,
                "periodic_tanh": 7, "relu": 8}

_d = ctypes.POINTER(ctypes.c_double)
_i = ctypes.POINTER(ctypes.c_int)
_ll = ctypes.POINTER(ctypes.c_longlong)
_vp = ctypes.c_void_p

# every symbol include/prophet_b200.h declares (tests check that the library exports exactly these)
SYMBOLS = [
    "pp_last_error", "pp_version", "pp_create", "pp_destroy", "pp_set_descriptors", "pp_num_descriptors",
    "pp_set_network", "pp_set_output_scaling", "pp_num_parameters", "pp_set_parameters", "pp_get_parameters",
    "pp_set_input_scaling", "pp_cart_to_frac", "pp_cart_to_frac_batch", "pp_add_structures", "pp_clear_structures", "pp_recompute_descriptors",
    "pp_num_structures",
    "pp_num_atoms", "pp_get_descriptors", "pp_get_neighbors", "pp_precondition", "pp_get_input_scaling", "pp_epoch",
    "pp_epoch_device", "pp_epoch_result_dev", "pp_epoch_form", "pp_evaluate", "pp_get_atom_energies", "pp_md_set_types",
    "pp_md_compute", "pp_md_build_neighbors", "pp_md_get_neighbors", "pp_md_step", "pp_md_step_device",
    "pp_md_exchange_setup", "pp_md_exchange_device", "pp_md_step_exchange",
    "pp_comm_unique_id", "pp_comm_init", "pp_comm_rank", "pp_comm_size", "pp_comm_path", "pp_kernel_launches", "pp_last_phase_ms",
    "pp_stream", "pp_column_sums_static", "pp_comm_allreduce_host",
]


class PPError(RuntimeError):
    pass


def lib():
    global _LIB
    if _LIB is None:
        if not os.path.exists(LIB_PATH):
            raise PPError("%s not built: run `python -c 'import __graft_entry__ as g; g.build()'` "
                          "(there is no CPU fallback)" % LIB_PATH)
        L = ctypes.CDLL(LIB_PATH)
        L.pp_last_error.restype = ctypes.c_char_p
        L.pp_num_atoms.restype = ctypes.c_longlong
        L.pp_kernel_launches.restype = ctypes.c_longlong
        L.pp_last_phase_ms.restype = ctypes.c_double
        L.pp_stream.restype = ctypes.c_void_p
        L.pp_set_output_scaling.argtypes = [_vp, ctypes.c_double, ctypes.c_double]
        L.pp_last_phase_ms.argtypes = [_vp, ctypes.c_int]
        L.pp_kernel_launches.argtypes = [_vp]
        L.pp_stream.argtypes = [_vp]
        L.pp_num_atoms.argtypes = [_vp]
        _LIB = L
    return _LIB


def _check(rc):
    if rc != 0:
        raise PPError(lib().pp_last_error().decode())


def _c(a, dtype):
    return np.ascontiguousarray(a, dtype=dtype)


def _p(a, t):
    return a.ctypes.data_as(t) if a is not None else None


def cart_to_frac(cell, cart):
    cell = _c(cell, np.float64); cart = _c(cart, np.float64)
    out = np.empty_like(cart)
    _check(lib().pp_cart_to_frac(_p(cell, _d), _p(cart, _d), ctypes.c_int(len(cart)), _p(out, _d)))
    return out


def cart_to_frac_batch(natoms, cells, cart):
    natoms = _c(natoms, np.int32); cells = _c(cells, np.float64); cart = _c(cart, np.float64)
    out = np.empty_like(cart)
    _check(lib().pp_cart_to_frac_batch(ctypes.c_int(len(natoms)), _p(natoms, _i), _p(cells, _d), _p(cart, _d), _p(out, _d)))
    return out


class Context:
    """One pp_context (one GPU).  Thin, argument-marshalling only."""

    def __init__(self, device: int = 0, fp_mode: int = 64):
        self._h = _vp()
        _check(lib().pp_create(ctypes.c_int(device), ctypes.c_int(fp_mode), ctypes.byref(self._h)))
        self.n_elem = 0

    @property
    def handle(self):
        """the pp_context* (for the host layer's C++ entry points that take a context)"""
        return self._h

    def close(self):
        if self._h:
            lib().pp_destroy(self._h)
            self._h = _vp()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    # ---- descriptors / networks ----
    def set_descriptors(self, n_elem, G1, G2, G3, G4, rcut):
        g1 = _c(np.asarray(G1, dtype=np.float64).reshape(-1, 2), np.float64)
        g2 = _c(np.asarray(G2, dtype=np.float64).reshape(-1, 4), np.float64)
        g3 = _c(np.asarray(G3, dtype=np.float64).reshape(-1, 5), np.float64)
        g4 = _c(np.asarray(G4, dtype=np.float64).reshape(-1, 5), np.float64)
        _check(lib().pp_set_descriptors(self._h, ctypes.c_int(n_elem), len(g1), _p(g1, _d), len(g2), _p(g2, _d),
                                        len(g3), _p(g3, _d), len(g4), _p(g4, _d), ctypes.c_double(rcut)))
        self.n_elem = n_elem

    @property
    def NG(self):
        return lib().pp_num_descriptors(self._h)

    def set_network(self, elem, sizes, params, mean, scale, transfer="tanh", node_transfer=None):
        sizes = _c(sizes, np.int32)
        n_nodes = int(sizes[1:].sum())
        if node_transfer is None:
            node_transfer = [transfer] * (n_nodes - 1) + ["linear"]
        tf = _c([TRANSFER_IDS[str(t)] for t in node_transfer], np.int32)
        params = _c(params, np.float64); mean = _c(mean, np.float64); scale = _c(scale, np.float64)
        _check(lib().pp_set_network(self._h, ctypes.c_int(elem), ctypes.c_int(len(sizes) - 1), _p(sizes, _i), _p(tf, _i),
                                    _p(params, _d), _p(mean, _d), _p(scale, _d)))

    def set_output_scaling(self, out_mean, out_scale):
        _check(lib().pp_set_output_scaling(self._h, out_mean, out_scale))

    @property
    def n_params(self):
        return lib().pp_num_parameters(self._h)

    def set_parameters(self, x):
        x = _c(x, np.float64)
        assert len(x) == self.n_params
        _check(lib().pp_set_parameters(self._h, _p(x, _d)))

    def get_parameters(self):
        x = np.zeros(self.n_params)
        _check(lib().pp_get_parameters(self._h, _p(x, _d)))
        return x

    def set_input_scaling(self, elem, mean, scale):
        mean = _c(mean, np.float64); scale = _c(scale, np.float64)
        _check(lib().pp_set_input_scaling(self._h, ctypes.c_int(elem), _p(mean, _d), _p(scale, _d)))

    def get_input_scaling(self, elem):
        NG = self.NG
        m = np.zeros(NG); s = np.zeros(NG)
        _check(lib().pp_get_input_scaling(self._h, ctypes.c_int(elem), _p(m, _d), _p(s, _d)))
        return m, s

    # ---- training set ----
    def add_structures(self, natoms, cells, frac, elem, targets, tags):
        natoms = _c(natoms, np.int32); cells = _c(cells, np.float64); frac = _c(frac, np.float64)
        elem = _c(elem, np.int32); targets = _c(targets, np.float64); tags = _c(tags, np.int32)
        assert cells.size == 9 * len(natoms) and frac.size == 3 * int(natoms.sum()) and len(elem) == int(natoms.sum())
        _check(lib().pp_add_structures(self._h, ctypes.c_int(len(natoms)), _p(natoms, _i), _p(cells, _d), _p(frac, _d),
                                       _p(elem, _i), _p(targets, _d), _p(tags, _i)))
        self._natoms = np.concatenate([getattr(self, "_natoms", np.zeros(0, dtype=np.int32)), natoms])

    def clear_structures(self):
        _check(lib().pp_clear_structures(self._h))
        self._natoms = np.zeros(0, dtype=np.int32)

    def recompute_descriptors(self):
        """init_G again after set_descriptors changed row parameters in place (Potential.cpp:806-808)."""
        _check(lib().pp_recompute_descriptors(self._h))

    @property
    def n_structures(self):
        return lib().pp_num_structures(self._h)

    @property
    def n_atoms(self):
        return lib().pp_num_atoms(self._h)

    def get_descriptors(self, s):
        out = np.zeros((self.NG, int(self._natoms[s])))
        _check(lib().pp_get_descriptors(self._h, ctypes.c_int(s), _p(out, _d)))
        return out

    def get_neighbors(self, s):
        n = int(self._natoms[s])
        npairs = ctypes.c_int(); nimg = ctypes.c_int()
        _check(lib().pp_get_neighbors(self._h, ctypes.c_int(s), None, None, None, None, ctypes.byref(npairs), ctypes.byref(nimg)))
        off = np.zeros(n + 1, dtype=np.int32); nj = np.zeros(max(npairs.value, 1), dtype=np.int32)
        ioff = np.zeros(npairs.value + 1, dtype=np.int32); it = np.zeros(max(nimg.value, 1), dtype=np.int32)
        _check(lib().pp_get_neighbors(self._h, ctypes.c_int(s), _p(off, _i), _p(nj, _i), _p(ioff, _i), _p(it, _i),
                                      ctypes.byref(npairs), ctypes.byref(nimg)))
        return dict(neigh_offset=off, neigh_j=nj[:npairs.value], image_offset=ioff, image_t=it[:nimg.value])

    def precondition(self, input_precondition=True):
        _check(lib().pp_precondition(self._h, ctypes.c_int(1 if input_precondition else 0)))

    # ---- epoch ----
    def epoch(self, params=None):
        P = self.n_params
        grad = np.zeros(P)
        sse = ctypes.c_double(); vsse = ctypes.c_double(); nval = ctypes.c_longlong()
        x = _c(params, np.float64) if params is not None else None
        _check(lib().pp_epoch(self._h, _p(x, _d), _p(grad, _d), ctypes.byref(sse), ctypes.byref(vsse), ctypes.byref(nval)))
        return dict(gradient=grad, SSE=sse.value, vSSE=vsse.value, nval=nval.value)

    def epoch_into(self, params, grad_out):
        """pp_epoch with caller-owned (e.g. pinned) host buffers; returns (sse, vsse, nval)."""
        sse = ctypes.c_double(); vsse = ctypes.c_double(); nval = ctypes.c_longlong()
        _check(lib().pp_epoch(self._h, _vp(params) if isinstance(params, int) else _p(params, _d),
                              _vp(grad_out) if isinstance(grad_out, int) else _p(grad_out, _d),
                              ctypes.byref(sse), ctypes.byref(vsse), ctypes.byref(nval)))
        return sse.value, vsse.value, nval.value

    def epoch_device(self):
        _check(lib().pp_epoch_device(self._h))

    def epoch_form(self):
        """0 two-phase with recompute, 1 fused (one CTA per SM), 2 fused with in-place deltas (two CTAs per SM), 3 two-phase
        with the activations stored in HBM between the phases."""
        return lib().pp_epoch_form(self._h)

    def epoch_result_dev(self):
        ptr = _d(); n = ctypes.c_int()
        _check(lib().pp_epoch_result_dev(self._h, ctypes.byref(ptr), ctypes.byref(n)))
        return ctypes.cast(ptr, ctypes.c_void_p).value, n.value

    def evaluate(self):
        E = np.zeros(self.n_structures)
        _check(lib().pp_evaluate(self._h, _p(E, _d)))
        return E

    def atom_energies(self, s):
        out = np.zeros(int(self._natoms[s]))
        _check(lib().pp_get_atom_energies(self._h, ctypes.c_int(s), _p(out, _d)))
        return out

    # ---- MD ----
    def md_set_types(self, type_to_elem: Sequence[int]):
        t = _c(type_to_elem, np.int32)
        _check(lib().pp_md_set_types(self._h, ctypes.c_int(len(t) - 1), _p(t, _i)))

    def md_compute(self, x, types, nlocal, offsets, neigh, f=None, want_G=False, want_dEdG=False):
        x = _c(x, np.float64); types = _c(types, np.int32); offsets = _c(offsets, np.int64); neigh = _c(neigh, np.int32)
        ntotal = len(x)
        if f is None:
            f = np.zeros((ntotal, 3))
        eatom = np.zeros(max(nlocal, 1)); et = ctypes.c_double()
        NG = self.NG
        G = np.zeros((nlocal, NG)) if want_G else None
        dEdG = np.zeros((nlocal, NG)) if want_dEdG else None
        _check(lib().pp_md_compute(self._h, _p(x, _d), _p(types, _i), ctypes.c_int(nlocal), ctypes.c_int(ntotal - nlocal),
                                   _p(offsets, _ll), _p(neigh, _i), _p(f, _d), _p(eatom, _d), ctypes.byref(et), _p(G, _d), _p(dEdG, _d)))
        return dict(f=f, eatom=eatom[:nlocal], etotal=et.value, G=G, dEdG=dEdG)

    def md_build_neighbors(self, x, types, nlocal, cutneigh):
        x = _c(x, np.float64); types = _c(types, np.int32)
        _check(lib().pp_md_build_neighbors(self._h, _p(x, _d), _p(types, _i), ctypes.c_int(nlocal), ctypes.c_int(len(x) - nlocal),
                                           ctypes.c_double(cutneigh)))
        self._md_nlocal = nlocal
        self._md_ntotal = len(x)

    def md_get_neighbors(self):
        total = ctypes.c_longlong()
        _check(lib().pp_md_get_neighbors(self._h, None, None, ctypes.byref(total)))
        off = np.zeros(self._md_nlocal + 1, dtype=np.int64); flat = np.zeros(max(total.value, 1), dtype=np.int32)
        _check(lib().pp_md_get_neighbors(self._h, _p(off, _ll), _p(flat, _i), ctypes.byref(total)))
        return off, flat[:total.value]

    def md_step(self, x, f=None):
        x = _c(x, np.float64)
        if f is None:
            f = np.zeros_like(x)
        eatom = np.zeros(max(self._md_nlocal, 1)); et = ctypes.c_double()
        _check(lib().pp_md_step(self._h, _p(x, _d), _p(f, _d), _p(eatom, _d), ctypes.byref(et)))
        return dict(f=f, eatom=eatom[:self._md_nlocal], etotal=et.value)

    def md_step_ptr(self, x_ptr: int, f_ptr: int, eatom_ptr: Optional[int] = None):
        """pp_md_step with raw host pointers (pinned buffers owned by the caller)."""
        et = ctypes.c_double()
        _check(lib().pp_md_step(self._h, _vp(x_ptr), _vp(f_ptr), _vp(eatom_ptr) if eatom_ptr else None, ctypes.byref(et)))
        return et.value

    def md_step_device(self, x_dev: int, f_dev: int, etotal_dev: int):
        _check(lib().pp_md_step_device(self._h, _vp(x_dev), _vp(f_dev), _vp(etotal_dev)))

    def md_exchange_setup(self, self_src, self_dst, peers):
        """peers: list of (rank, send_idx array, recv_dst array)."""
        self_src = _c(self_src, np.int32); self_dst = _c(self_dst, np.int32)
        ranks = _c([p[0] for p in peers], np.int32)
        sc = _c([len(p[1]) for p in peers], np.int32); rc = _c([len(p[2]) for p in peers], np.int32)
        sflat = _c(np.concatenate([np.asarray(p[1], dtype=np.int32) for p in peers]) if peers else np.zeros(0), np.int32)
        rflat = _c(np.concatenate([np.asarray(p[2], dtype=np.int32) for p in peers]) if peers else np.zeros(0), np.int32)
        _check(lib().pp_md_exchange_setup(self._h, ctypes.c_int(len(self_src)), _p(self_src, _i), _p(self_dst, _i), ctypes.c_int(len(peers)),
                                          _p(ranks, _i), _p(sc, _i), _p(sflat, _i), _p(rc, _i), _p(rflat, _i)))

    def md_exchange_device(self, f_dev: int, etotal_dev: int = 0):
        _check(lib().pp_md_exchange_device(self._h, _vp(f_dev), _vp(etotal_dev) if etotal_dev else None))

    def md_step_exchange_ptr(self, x_ptr: int, f_ptr: int, eatom_ptr: Optional[int] = None):
        et = ctypes.c_double()
        _check(lib().pp_md_step_exchange(self._h, _vp(x_ptr), _vp(f_ptr), _vp(eatom_ptr) if eatom_ptr else None, ctypes.byref(et)))
        return et.value

    # ---- comm / instrumentation ----
    @staticmethod
    def comm_unique_id() -> bytes:
        buf = ctypes.create_string_buffer(128)
        _check(lib().pp_comm_unique_id(buf))
        return buf.raw

    def comm_init(self, nranks, rank, unique_id: bytes):
        buf = ctypes.create_string_buffer(unique_id, 128)
        _check(lib().pp_comm_init(self._h, ctypes.c_int(nranks), ctypes.c_int(rank), buf))

    def comm_path(self) -> str:
        lib().pp_comm_path.restype = ctypes.c_char_p
        return lib().pp_comm_path(self._h).decode()

    @property
    def kernel_launches(self):
        return lib().pp_kernel_launches(self._h)

    def phase_ms(self, which):
        return lib().pp_last_phase_ms(self._h, which)

    @property
    def stream(self):
        return lib().pp_stream(self._h)
