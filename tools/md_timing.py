#!/usr/bin/env python
"""pair_nn step timing on one GPU: a diamond-Si box of cells^3 unit cells (C2's rows and network), neighbour list built on
the GPU, then `steps` calls of pp_md_step with the library's own phase timers (descriptors / network / forces).  Also prints a
checksum of the forces so two builds (PP_LIB_PATH=...) can be compared bit for bit.
Usage: python tools/md_timing.py [cells=30] [steps=3]"""
import hashlib
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
from prophet_b200 import capi, synth  # noqa: E402

cells = int(sys.argv[1]) if len(sys.argv) > 1 else 30
steps = int(sys.argv[2]) if len(sys.argv) > 2 else 3
cfg = synth.config_c5(cells)
cell, cart, Z, _, _ = synth.make_structure(cfg, 5)
x, _, nlocal = synth.expand_ghosts(cell, cart, Z, cfg.rcut + 1.0)
x = np.ascontiguousarray(x)
ctx = capi.Context(0)
ctx.set_descriptors(1, cfg.G1, cfg.G2, cfg.G3, cfg.G4, cfg.rcut)
net = synth.random_network(cfg, 14, 1)
mean = 0.1 * synth.normal(9, cfg.NG, 3)
scale = 0.5 + np.abs(synth.normal(9, cfg.NG, 4))
ctx.set_network(0, net["sizes"], net["params"], mean, scale)
ctx.md_set_types([0, 0])
ctx.md_build_neighbors(x, np.ones(len(x), dtype=np.int32), nlocal, cfg.rcut + 1.0)
f = np.zeros_like(x)
best = None
for s in range(steps + 1):
    f[:] = 0.0
    t0 = time.perf_counter()
    r = ctx.md_step(x, f)
    wall = time.perf_counter() - t0
    ph = [ctx.phase_ms(k) for k in (4, 5, 6)]
    if s > 0 and (best is None or sum(ph) < sum(best)):
        best = ph
print("atoms %d + %d ghosts: descriptors %.3f ms, network %.3f ms, forces %.3f ms, sum %.3f ms (best of %d); last wall %.1f ms"
      % (nlocal, len(x) - nlocal, best[0], best[1], best[2], sum(best), steps, wall * 1e3))
print("etotal %.15e  |f| %.15e  sha1(f) %s" % (r["etotal"], float(np.linalg.norm(f)), hashlib.sha1(f.tobytes()).hexdigest()[:16]))
ctx.close()
