#!/usr/bin/env python
"""One process for an `ncu --set full` capture of the kernels that changed late in round 2: two epochs of C3 (2 000
structures: mlp_fused2_kernel<FORWARD> / <TRAIN> with stored activations, epoch form 5) and two MD steps of a 216 000-atom
box (md_desc_kernel, md_force_kernel).  Usage: ncu ... --kernel-name regex:"mlp_fused2_kernel|md_(desc|force)_kernel" python tools/profile_case.py"""
import os
import sys

import numpy as np

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench  # noqa: E402
from prophet_b200 import capi, synth  # noqa: E402

cfg = synth.CONFIGS["C3"](2000)
ctx, nets = bench.make_context(capi, synth, cfg, 0)
bench.load_shard(ctx, capi, synth, cfg, np.arange(2000))
ctx.precondition(True)
x0 = np.concatenate([n["params"] for n in nets])
for _ in range(2):
    r = ctx.epoch(x0)
print("C3 2000: form", ctx.epoch_form(), "SSE %.6e" % r["SSE"], flush=True)
ctx.close()

cfg = synth.config_c5(30)
cell, cart, Z, _, _ = synth.make_structure(cfg, 5)
x, _, nlocal = synth.expand_ghosts(cell, cart, Z, cfg.rcut + 1.0)
x = np.ascontiguousarray(x)
ctx = capi.Context(0)
ctx.set_descriptors(1, cfg.G1, cfg.G2, cfg.G3, cfg.G4, cfg.rcut)
net = synth.random_network(cfg, 14, 1)
ctx.set_network(0, net["sizes"], net["params"], np.zeros(cfg.NG), np.ones(cfg.NG))
ctx.md_set_types([0, 0])
ctx.md_build_neighbors(x, np.ones(len(x), dtype=np.int32), nlocal, cfg.rcut + 1.0)
for _ in range(2):
    r = ctx.md_step(x)
print("MD 216000: etotal %.9e" % r["etotal"], flush=True)
ctx.close()
