"""ctypes access to the host-side C++ layer (libprophet_b200_host.so): test hooks for Setup, the checkpoint
reader/writer, the Optimizer and the pair_style nn shim.  Harness code; the classes themselves are C++."""
import ctypes
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libprophet_b200_host.so")
CLI_PATH = os.path.join(_HERE, "PROPhet_b200")
_LIB = None
_d = ctypes.POINTER(ctypes.c_double)
_i = ctypes.POINTER(ctypes.c_int)


def lib():
    global _LIB
    if _LIB is None:
        if not os.path.exists(LIB_PATH):
            raise RuntimeError("%s not built (make -C prophet_b200/host)" % LIB_PATH)
        # the device library first, by full path, so the host library's dependency resolves to the in-tree build
        ctypes.CDLL(os.path.join(_HERE, "libprophet_b200.so"), mode=ctypes.RTLD_GLOBAL)
        L = ctypes.CDLL(LIB_PATH)
        L.pph_optimizer_create.restype = ctypes.c_void_p
        L.pph_optimizer_create.argtypes = [ctypes.c_int, _d, ctypes.c_char_p, ctypes.c_int]
        L.pph_optimizer_create2.restype = ctypes.c_void_p
        L.pph_optimizer_create2.argtypes = [ctypes.c_int, _d, ctypes.c_char_p, ctypes.c_int, ctypes.c_double,
                                            ctypes.c_double, ctypes.c_double]
        L.pph_optimizer_step.argtypes = [ctypes.c_void_p, ctypes.c_int, ctypes.c_double, _d, _d]
        L.pph_optimizer_step_ptr.argtypes = [ctypes.c_void_p, ctypes.c_int, ctypes.c_double, ctypes.c_void_p, ctypes.c_void_p]
        L.pph_optimizer_destroy.argtypes = [ctypes.c_void_p]
        _LIB = L
    return _LIB


class Optimizer:
    """pph::Optimizer (Optimizer::update_network + rprop / steepest descent) driven with caller-supplied gradients."""

    def __init__(self, x0, algorithm="rprop", nsystems=1, line_min_epsilon=None, debug=None, sd_momentum=None):
        x0 = np.ascontiguousarray(x0, dtype=np.float64)
        self.P = len(x0)
        if line_min_epsilon is None and debug is None and sd_momentum is None:
            self._h = lib().pph_optimizer_create(self.P, x0.ctypes.data_as(_d), algorithm.encode(), nsystems)
        else:     # Functional_params.cpp:56-103 defaults for what is not given
            self._h = lib().pph_optimizer_create2(self.P, x0.ctypes.data_as(_d), algorithm.encode(), nsystems,
                                                  line_min_epsilon or 0.0, debug or 0.0,
                                                  sd_momentum or 0.0)

    def configure(self, x0, lambda_max=0.0, ncycles=1, threshold=0.0, do_print=False):
        """Re-create the optimizer with lambda_max / Ncycles / threshold (input_file keys) and printing on or off."""
        x0 = np.ascontiguousarray(x0, dtype=np.float64)
        L = lib()
        L.pph_optimizer_configure.argtypes = [ctypes.c_void_p, ctypes.c_double, ctypes.c_int, ctypes.c_double, ctypes.c_int,
                                              ctypes.c_int, _d]
        L.pph_optimizer_converged.argtypes = [ctypes.c_void_p]
        L.pph_optimizer_alpha.argtypes = [ctypes.c_void_p]
        L.pph_optimizer_alpha.restype = ctypes.c_double
        L.pph_optimizer_configure(self._h, lambda_max, ncycles, threshold, int(do_print), self.P, x0.ctypes.data_as(_d))

    def converged(self):
        return bool(lib().pph_optimizer_converged(self._h))

    def alpha(self):
        return lib().pph_optimizer_alpha(self._h)

    def step(self, sse, grad):
        grad = np.ascontiguousarray(grad, dtype=np.float64)
        x = np.zeros(self.P)
        lib().pph_optimizer_step(self._h, self.P, float(sse), grad.ctypes.data_as(_d), x.ctypes.data_as(_d))
        return x

    def train_loop(self, ctx_handle, x_ptr, grad_ptr, steps):
        """`steps` iterations of the host layer's training loop (pp_epoch + Optimizer::update_network) in C++;
        returns (seconds, last SSE).  x_ptr / grad_ptr: host buffers of P doubles (parameters in/out, gradient scratch)."""
        L = lib()
        L.pph_train_loop.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_int, ctypes.c_void_p, ctypes.c_void_p,
                                     ctypes.c_int, ctypes.POINTER(ctypes.c_double), ctypes.POINTER(ctypes.c_double)]
        t = ctypes.c_double(); sse = ctypes.c_double()
        rc = L.pph_train_loop(ctx_handle, self._h, self.P, x_ptr, grad_ptr, steps, ctypes.byref(t), ctypes.byref(sse))
        if rc:
            raise RuntimeError("pph_train_loop failed")
        return t.value, sse.value

    def step_ptr(self, x_ptr, grad_ptr, sse=0.0):
        lib().pph_optimizer_step_ptr(self._h, self.P, float(sse), grad_ptr, x_ptr)

    def __del__(self):
        try:
            lib().pph_optimizer_destroy(self._h)
        except Exception:
            pass


class Rprop(Optimizer):
    def __init__(self, P_or_x0):
        x0 = np.zeros(P_or_x0) if np.isscalar(P_or_x0) else P_or_x0
        super().__init__(x0, "rprop", 1)


def ks(a, b):
    """approximate Kolmogorov-Smirnov statistic of Potential::validate -> (D, p)."""
    a = np.ascontiguousarray(a, dtype=np.float64)
    b = np.ascontiguousarray(b, dtype=np.float64)
    p = ctypes.c_double(0.0)
    L = lib()
    L.pph_ks.restype = ctypes.c_double
    L.pph_ks.argtypes = [_d, ctypes.c_int, _d, ctypes.c_int, ctypes.POINTER(ctypes.c_double)]
    D = L.pph_ks(a.ctypes.data_as(_d), len(a), b.ctypes.data_as(_d), len(b), ctypes.byref(p))
    return D, p.value


def checkpoint_roundtrip(in_path, out_path):
    rc = lib().pph_checkpoint_roundtrip(in_path.encode(), out_path.encode())
    if rc:
        raise RuntimeError("could not read " + in_path)


def setup_parse(input_file, rank=0, nprocs=1):
    buf = ctypes.create_string_buffer(1 << 20)
    lib().pph_setup_parse(input_file.encode(), rank, nprocs, buf, len(buf))
    return buf.value.decode()


def read_vasprun(directory, cap=4096):
    cell = np.zeros(9); pos = np.zeros(3 * cap); Z = np.zeros(cap, dtype=np.int32)
    n = ctypes.c_int(); tgt = ctypes.c_double()
    rc = lib().pph_read_vasprun(directory.encode(), cell.ctypes.data_as(_d), ctypes.byref(n), pos.ctypes.data_as(_d),
                                Z.ctypes.data_as(_i), ctypes.byref(tgt), cap)
    assert rc == 0
    return cell.reshape(3, 3), pos[:3 * n.value].reshape(-1, 3), Z[:n.value].copy(), tgt.value


def auto_rows(rcut, nradial, nangular, cap=256):
    g2 = np.zeros(4 * cap); g4 = np.zeros(5 * cap); n2 = ctypes.c_int(); n4 = ctypes.c_int()
    rc = lib().pph_auto_rows(ctypes.c_double(rcut), nradial, nangular, g2.ctypes.data_as(_d), ctypes.byref(n2),
                             g4.ctypes.data_as(_d), ctypes.byref(n4), cap)
    assert rc == 0
    return g2[:4 * n2.value].reshape(-1, 4), g4[:5 * n4.value].reshape(-1, 5)


def pair_nn_run(rcut, coeff_types, coeff_files, x, types, nlocal, numneigh, neigh_flat):
    """pair_style nn <rcut>; pair_coeff <type> <file> ...; one compute() through the LAMMPS stand-in."""
    x = np.ascontiguousarray(x, dtype=np.float64); types = np.ascontiguousarray(types, dtype=np.int32)
    numneigh = np.ascontiguousarray(numneigh, dtype=np.int32); neigh_flat = np.ascontiguousarray(neigh_flat, dtype=np.int32)
    ct = np.ascontiguousarray(coeff_types, dtype=np.int32)
    files = (ctypes.c_char_p * len(coeff_files))(*[f.encode() for f in coeff_files])
    f = np.zeros_like(x); eatom = np.zeros(nlocal); eng = ctypes.c_double(); restart_ok = ctypes.c_int(0)
    rc = lib().pph_pair_nn_run(ctypes.c_double(rcut), len(ct), ct.ctypes.data_as(_i), files, nlocal, len(x) - nlocal,
                               x.ctypes.data_as(_d), types.ctypes.data_as(_i), numneigh.ctypes.data_as(_i),
                               neigh_flat.ctypes.data_as(_i), f.ctypes.data_as(_d), eatom.ctypes.data_as(_d), ctypes.byref(eng),
                               ctypes.byref(restart_ok))
    assert rc == 0, rc      # 2: init_style did not request the reference's neighbour settings, 3: single(), 4: no temp file
    return dict(f=f, eatom=eatom, eng_vdwl=eng.value, restart_ok=bool(restart_ok.value))
