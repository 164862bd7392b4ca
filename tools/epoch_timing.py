#!/usr/bin/env python
"""Epoch time of any synthetic configuration (C2 / C3 / C4) at a chosen structure count (tuning aid).
    python tools/epoch_timing.py C3 2000 [transfer]        PP_FP_MODE=32 selects the FP32 mode of pp_create"""
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
from prophet_b200 import capi, synth  # noqa: E402


def main():
    name = sys.argv[1] if len(sys.argv) > 1 else "C3"
    n_struct = int(sys.argv[2]) if len(sys.argv) > 2 else 2000
    cfg = synth.CONFIGS[name](n_struct)
    if len(sys.argv) > 3:
        cfg.transfer = sys.argv[3]          # e.g. tanh_spline (the reference's default transfer function)
    ctx = capi.Context(0, int(os.environ.get("PP_FP_MODE", "64")))
    ctx.set_descriptors(len(cfg.elements), cfg.G1, cfg.G2, cfg.G3, cfg.G4, cfg.rcut)
    zmap = {z: i for i, z in enumerate(cfg.elements)}
    for e, z in enumerate(cfg.elements):
        net = synth.random_network(cfg, z, 1)
        n_nodes = int(sum(net["sizes"][1:]))
        ctx.set_network(e, net["sizes"], net["params"], np.zeros(cfg.NG), np.ones(cfg.NG),
                        node_transfer=[cfg.transfer] * (n_nodes - 1) + ["linear"])
    n = cfg.natoms
    desc_ms = 0.0
    for first in range(0, n_struct, 500):
        cnt = min(500, n_struct - first)
        b = synth.make_batch(cfg, first, cnt)
        frac = np.empty_like(b["cart"])
        for k in range(cnt):
            frac[k * n:(k + 1) * n] = capi.cart_to_frac(b["cells"][k], b["cart"][k * n:(k + 1) * n])
        elem = np.array([zmap[int(z)] for z in b["Z"]], dtype=np.int32)
        ctx.add_structures(b["natoms"], b["cells"], frac, elem, b["targets"], b["tags"])
        desc_ms += ctx.phase_ms(0)
    ctx.precondition(True)
    stream = torch.cuda.ExternalStream(ctx.stream)
    for _ in range(3):
        ctx.epoch_device()
    e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
    reps = 10
    e0.record(stream)
    for _ in range(reps):
        ctx.epoch_device()
    e1.record(stream)
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / reps
    sizes = [cfg.NG] + list(cfg.hidden) + [1]
    F = 2 * sum(a * b for a, b in zip(sizes[:-1], sizes[1:]))
    flops = (3 * F - 2 * sizes[0] * sizes[1]) * n_struct * n
    ctx.epoch()
    print("%s: %d structures x %d atoms, NG %d, nets %s, P %d: epoch %.3f ms (form %d), %.2f M structures/s, %.2f TFLOP/s algorithmic; "
          "descriptors %.1f ms (%.0f structures/s)" % (name, n_struct, n, cfg.NG, sizes, ctx.n_params, ms, ctx.epoch_form(),
                                                      n_struct / ms / 1e3, flops / ms / 1e9, desc_ms, n_struct / desc_ms * 1e3))
    print("   phases ms: forward %.3f backward+reduce %.3f collective %.3f" % (ctx.phase_ms(1), ctx.phase_ms(2), ctx.phase_ms(3)))
    ctx.close()


if __name__ == "__main__":
    main()
