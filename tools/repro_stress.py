#!/usr/bin/env python
"""Bit-reproducibility stress of the training epoch (a shared-memory race would show up as run-to-run differences)."""
import os, sys
import numpy as np
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
from prophet_b200 import capi, synth  # noqa: E402

n_struct = int(sys.argv[1]) if len(sys.argv) > 1 else 4000
reps = int(sys.argv[2]) if len(sys.argv) > 2 else 200
cfg = synth.config_c2(n_struct)
ctx = capi.Context(0)
ctx.set_descriptors(1, cfg.G1, cfg.G2, cfg.G3, cfg.G4, cfg.rcut)
net = synth.random_network(cfg, 14, 1)
ctx.set_network(0, net["sizes"], net["params"], np.zeros(cfg.NG), np.ones(cfg.NG))
n = cfg.natoms
for first in range(0, n_struct, 1000):
    cnt = min(1000, n_struct - first)
    b = synth.make_batch(cfg, first, cnt)
    frac = np.empty_like(b["cart"])
    for k in range(cnt):
        frac[k * n:(k + 1) * n] = capi.cart_to_frac(b["cells"][k], b["cart"][k * n:(k + 1) * n])
    ctx.add_structures(b["natoms"], b["cells"], frac, np.zeros(cnt * n, dtype=np.int32), b["targets"], b["tags"])
ctx.precondition(True)
for env in (None, "1"):
    if env: os.environ["PP_MLP_ONE_CTA"] = env
    else: os.environ.pop("PP_MLP_ONE_CTA", None)
    ref = ctx.epoch(net["params"])
    bad = 0
    for _ in range(reps):
        r = ctx.epoch(net["params"])
        if not (np.array_equal(r["gradient"], ref["gradient"]) and r["SSE"] == ref["SSE"] and r["vSSE"] == ref["vSSE"]):
            bad += 1
    print("form %d: %d epochs, %d differ from the first" % (ctx.epoch_form(), reps, bad))
ctx.close()
