#!/usr/bin/env python
"""Per-phase cycle counts of the fused training kernel (tuning aid, not part of the product or the bench).

Needs a library built with the timing hooks:  make -C prophet_b200/csrc PTXAS=-DPP_MLP_TIMING  (rebuild without
the flag afterwards).  Prints, for warp 0 and the last warp of CTA 0, the average cycles per tile spent between
consecutive PP_T() stamps of mlp_kernel (see pp_mlp.cu)."""
import ctypes
import os
import sys

import numpy as np

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
from prophet_b200 import capi, synth  # noqa: E402

NAMES = ["wait tile + sync", "issue next load + struct table", "fwd gemm (sum over layers)", "fwd transfer",
         "fwd pair barrier", "output node", "sync before error", "per-structure error", "delta scale + barrier",
         "bwd gemm + multiply", "bwd pair barrier", "sync before dW", "dW output node", "dW tiles", "fwd accumulator drain"]


def main():
    n_struct = int(sys.argv[1]) if len(sys.argv) > 1 else 148 * 16
    ctx = capi.Context(0)
    cfg = synth.config_c2(n_struct)
    ctx.set_descriptors(1, cfg.G1, cfg.G2, cfg.G3, cfg.G4, cfg.rcut)
    net = synth.random_network(cfg, 14, 1)
    ctx.set_network(0, net["sizes"], net["params"], np.zeros(cfg.NG), np.ones(cfg.NG))
    for first in range(0, n_struct, 592):
        cnt = min(592, n_struct - first)
        b = synth.make_batch(cfg, first, cnt)
        n = cfg.natoms
        frac = np.empty_like(b["cart"])
        for k in range(cnt):
            frac[k * n:(k + 1) * n] = capi.cart_to_frac(b["cells"][k], b["cart"][k * n:(k + 1) * n])
        ctx.add_structures(b["natoms"], b["cells"], frac, np.zeros(cnt * n, dtype=np.int32), b["targets"], b["tags"])
    ctx.precondition(True)
    L = capi.lib()
    L.pp_debug_mlp_cycles.argtypes = [ctypes.c_void_p, ctypes.c_int]
    for _ in range(3):
        ctx.epoch_device()
    ctx.epoch()      # synchronising call
    buf = (ctypes.c_ulonglong * 32)()
    L.pp_debug_mlp_cycles(buf, 1)
    reps = 10
    for _ in range(reps):
        ctx.epoch_device()
    ctx.epoch()      # synchronising call (the context's stream does not synchronise with the symbol copy)
    reps += 1
    L.pp_debug_mlp_cycles(buf, 0)
    c = np.array(list(buf), dtype=np.float64).reshape(2, 16)
    tiles = reps * n_struct / 148.0
    print("structures %d, tiles per CTA per epoch %.1f" % (n_struct, n_struct / 148.0))
    for w, name in ((0, "warp 0"), (1, "last warp")):
        print("   whole tile loop: %.0f cycles per tile" % (c[w][15] / tiles))
        c[w][15] = 0
        tot = c[w].sum() / tiles
        print("%s: %.0f cycles per tile" % (name, tot))
        for i, nm in enumerate(NAMES):
            print("   %-34s %8.0f  %5.1f%%" % (nm, c[w][i] / tiles, 100 * c[w][i] / max(c[w].sum(), 1)))


if __name__ == "__main__":
    main()
