#!/usr/bin/env python
"""Small cases for `compute-sanitizer --tool racecheck|memcheck|synccheck`: one training epoch in each kernel form that
ships (C2: two CTAs per SM; C3 shape: two-phase with recompute; C4 shape: stored two-phase; FP32 mode) and one MD step
(descriptors -> network -> forces) on a small periodic box.  Prints a checksum per case so the plain run can be compared.
Usage: compute-sanitizer --tool racecheck python tools/sanitize_case.py [c2|c3|c4|fp32|md ...]"""
import os
import sys

import numpy as np

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
from prophet_b200 import capi, synth  # noqa: E402


def train_case(name, n_struct, fp_mode=64):
    cfg = synth.CONFIGS[name](n_struct)
    ctx = capi.Context(0, fp_mode) if fp_mode != 64 else capi.Context(0)
    ctx.set_descriptors(len(cfg.elements), cfg.G1, cfg.G2, cfg.G3, cfg.G4, cfg.rcut)
    params = []
    for e, z in enumerate(cfg.elements):
        net = synth.random_network(cfg, z, 1)
        ctx.set_network(e, net["sizes"], net["params"], np.zeros(cfg.NG), np.ones(cfg.NG))
        params.append(net["params"])
    lut = np.zeros(128, dtype=np.int32)
    for i, z in enumerate(cfg.elements):
        lut[z] = i
    b = synth.make_batch(cfg, 0, n_struct)
    frac = capi.cart_to_frac_batch(b["natoms"], b["cells"], b["cart"])
    ctx.add_structures(b["natoms"], b["cells"], frac, lut[b["Z"]], b["targets"], b["tags"])
    ctx.precondition(True)
    r = ctx.epoch(np.concatenate(params))
    print("%s fp%d: form %d SSE %.12e |grad| %.12e" % (name, fp_mode, ctx.epoch_form(), r["SSE"], float(np.linalg.norm(r["gradient"]))), flush=True)
    ctx.close()


def md_case(cells=3):
    cfg = synth.config_c5(cells)
    cell, cart, Z, _, _ = synth.make_structure(cfg, 5)
    x, _, nlocal = synth.expand_ghosts(cell, cart, Z, cfg.rcut + 1.0)
    x = np.ascontiguousarray(x)
    ctx = capi.Context(0)
    ctx.set_descriptors(1, cfg.G1, cfg.G2, cfg.G3, cfg.G4, cfg.rcut)
    net = synth.random_network(cfg, 14, 1)
    ctx.set_network(0, net["sizes"], net["params"], np.zeros(cfg.NG), np.ones(cfg.NG))
    ctx.md_set_types([0, 0])
    ctx.md_build_neighbors(x, np.ones(len(x), dtype=np.int32), nlocal, cfg.rcut + 1.0)
    r = ctx.md_step(x)
    print("md: %d local + %d ghost atoms, etotal %.12e |f| %.12e" % (nlocal, len(x) - nlocal, r["etotal"], float(np.linalg.norm(r["f"]))), flush=True)
    ctx.close()


if __name__ == "__main__":
    which = sys.argv[1:] or ["c2", "c3", "c4", "fp32", "md"]
    if "c2" in which:
        train_case("C2", 24)
    if "c3" in which:
        train_case("C3", 8)
    if "c4" in which:
        train_case("C4", 3)
    if "fp32" in which:
        train_case("C2", 24, 32)
    if "md" in which:
        md_case()
