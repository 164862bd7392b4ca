#!/usr/bin/env python
"""Per-phase cycle counts of the stored two-phase kernels (tuning aid, not part of the product or the bench).

Needs a library whose pp_mlpw.cu was compiled with -DPP_MLPW_TIMING, linked into tools/_timing/libprophet_b200.so
(the product library is left alone).  Prints, for warps 0 and 8 of CTA 0, the cycles per 64 rows between consecutive PW_T()
stamps of mlpw_forward_kernel / mlpw_backward_kernel (see pp_mlpw.cu).
    python tools/mlpw_phase_timing.py C4 1000"""
import ctypes
import os
import sys

import numpy as np

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
from prophet_b200 import capi, synth  # noqa: E402

_T = os.path.join(os.path.dirname(os.path.abspath(__file__)), "_timing", "libprophet_b200.so")
if os.path.exists(_T):
    capi.LIB_PATH = _T          # the instrumented build, kept beside the tool (tools/_timing/, not the product library)

FWD = ["wait X", "barrier after normalise", "products (all layers)", "transfer + stores", "layer barriers", "output node", "end barrier", "normalise"]
BWD = ["wait H/T' of last layer", "output node + delta", "barrier", "T' loads + delta product in place", "normalise X", "wait H_l",
       "weight-gradient blocks", "barrier after them", "-", "barrier after delta"]


def main():
    name = sys.argv[1] if len(sys.argv) > 1 else "C4"
    n_struct = int(sys.argv[2]) if len(sys.argv) > 2 else 1000
    cfg = synth.CONFIGS[name](n_struct)
    ctx = capi.Context(0)
    ctx.set_descriptors(len(cfg.elements), cfg.G1, cfg.G2, cfg.G3, cfg.G4, cfg.rcut)
    zmap = {z: i for i, z in enumerate(cfg.elements)}
    for e, z in enumerate(cfg.elements):
        net = synth.random_network(cfg, z, 1)
        ctx.set_network(e, net["sizes"], net["params"], np.zeros(cfg.NG), np.ones(cfg.NG))
    n = cfg.natoms
    for first in range(0, n_struct, 500):
        cnt = min(500, n_struct - first)
        b = synth.make_batch(cfg, first, cnt)
        frac = np.concatenate([capi.cart_to_frac(b["cells"][k], b["cart"][k * n:(k + 1) * n]) for k in range(cnt)])
        ctx.add_structures(b["natoms"], b["cells"], frac, np.array([zmap[int(z)] for z in b["Z"]], dtype=np.int32), b["targets"], b["tags"])
    ctx.precondition(True)
    L = capi.lib()
    L.pp_debug_mlpw_cycles.argtypes = [ctypes.c_void_p, ctypes.c_int]
    for _ in range(3):
        ctx.epoch_device()
    ctx.epoch()
    buf = (ctypes.c_ulonglong * 64)()
    L.pp_debug_mlpw_cycles(buf, 1)
    reps = 5
    for _ in range(reps):
        ctx.epoch()
    L.pp_debug_mlpw_cycles(buf, 0)
    c = np.array(list(buf), dtype=np.float64).reshape(2, 2, 16)
    rows64 = reps * n_struct * n / 148.0 / 64.0          # 64-row units per CTA over the timed epochs
    print("%s, %d structures, form %d: %.1f 64-row units per CTA per epoch" % (name, n_struct, ctx.epoch_form(), n_struct * n / 148.0 / 64))
    for k, (kn, names) in enumerate((("forward", FWD), ("backward", BWD))):
        for w in range(2):
            tot = c[k][w].sum() / rows64
            print("%s kernel, warp %d: %.0f cycles per 64 rows of the CTA" % (kn, 8 * w, tot))
            for i, nm in enumerate(names):
                print("   %-30s %8.0f  %5.1f%%" % (nm, c[k][w][i] / rows64, 100 * c[k][w][i] / max(c[k][w].sum(), 1)))
    ctx.close()


if __name__ == "__main__":
    main()
