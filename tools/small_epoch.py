#!/usr/bin/env python
"""A small C2 training epoch in both fused forms plus an MD evaluation (for compute-sanitizer runs)."""
import os, sys
import numpy as np
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
from prophet_b200 import capi, synth  # noqa: E402

cfg = synth.config_c2(300)
ctx = capi.Context(0)
ctx.set_descriptors(1, cfg.G1, cfg.G2, cfg.G3, cfg.G4, cfg.rcut)
net = synth.random_network(cfg, 14, 1)
ctx.set_network(0, net["sizes"], net["params"], np.zeros(cfg.NG), np.ones(cfg.NG))
b = synth.make_batch(cfg, 0, 300)
n = cfg.natoms
frac = np.empty_like(b["cart"])
for k in range(300):
    frac[k * n:(k + 1) * n] = capi.cart_to_frac(b["cells"][k], b["cart"][k * n:(k + 1) * n])
ctx.add_structures(b["natoms"], b["cells"], frac, np.zeros(300 * n, dtype=np.int32), b["targets"], b["tags"])
ctx.precondition(True)
r2 = ctx.epoch(net["params"]); f2 = ctx.epoch_form()
os.environ["PP_MLP_ONE_CTA"] = "1"
r1 = ctx.epoch(net["params"]); f1 = ctx.epoch_form()
print("forms", f2, f1, "SSE", r2["SSE"], r1["SSE"], "max grad diff", np.abs(r2["gradient"] - r1["gradient"]).max())
E = ctx.evaluate()
print("E[0]", E[0])
ctx.close()
