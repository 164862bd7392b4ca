#!/usr/bin/env python
"""Wall time per step of the host layer's C++ training loop on C2 (tuning aid): pp_epoch + Optimizer::update_network."""
import os, sys
import numpy as np
import torch
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
from prophet_b200 import capi, hostapi, synth  # noqa: E402

n_struct = int(sys.argv[1]) if len(sys.argv) > 1 else 10000
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
P = ctx.n_params
x = torch.empty(P, dtype=torch.float64).pin_memory(); g = torch.empty(P, dtype=torch.float64).pin_memory()
x.numpy()[:] = net["params"]
opt = hostapi.Optimizer(net["params"], "rprop", int(0.95 * n_struct))
opt.train_loop(ctx.handle, x.data_ptr(), g.data_ptr(), 5)
for rep in range(3):
    t, sse = opt.train_loop(ctx.handle, x.data_ptr(), g.data_ptr(), 40)
    print("rep %d: %.1f us per step, sse %.6g, form %d" % (rep, t / 40 * 1e6, sse, ctx.epoch_form()))
ctx.close()
