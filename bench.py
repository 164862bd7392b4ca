#!/usr/bin/env python
"""Benchmark of the PROPhet Behler-Parrinello hot path on B200 (see DESIGN.md section "Measurement").

    python bench.py --gpus N --steps K --warmup W            # this repository's CUDA path
    python bench.py --impl reference --gpus N --steps K ...   # the reference's own CPU code on the host cores

Headline metric (BASELINE.json): training structures/sec/epoch on config C2 (10 000 synthetic 64-atom Si
structures per GPU, 50 descriptors, net 50-40-40-1 tanh).  A "step" is one training epoch = the body of
Potential::train's loop (Potential.cpp:626-672): per-element MLP forward, per-structure error, backprop to all
parameters, gradient reduction (+ NCCL allreduce for N > 1).
  value : descriptors and parameters resident in HBM, device-timed (CUDA events on the library's stream)
  e2e   : the same epoch through the C-ABI call pp_epoch with pinned HOST buffers: parameters host->device and
          [gradient, SSE, vSSE, nval] device->host inside the timed region, followed by the host optimizer step
Inputs (256 MB of FP64 descriptors per GPU) are larger than the 126 MB L2, so no explicit flush is needed.
"""
import argparse
import json
import os
import re
import shutil
import subprocess
import sys
import tempfile
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
# stdout must carry exactly ONE JSON line, but libraries (NCCL's version banner, for one) write to file descriptor 1
# directly.  Keep a private copy of the real stdout for the result and point fd 1 at stderr for everything else.
_RESULT_FD = os.dup(1)
os.dup2(2, 1)


def emit(obj):
    os.write(_RESULT_FD, (json.dumps(obj) + "\n").encode())

METRIC = "training structures/sec/epoch"
UNIT = "structures/s"


# ------------------------------------------------------------------------------------------------------------
class ClockSampler:
    """nvidia-smi clock / throttle-reason samples during the timed region (B200_PROFILING.md)."""

    def __init__(self, gpu_index=0):
        self.rows = []
        self.gpu = gpu_index
        self.proc = None

    def start(self):
        q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
             "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
             "clocks_event_reasons.sw_power_cap")
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.gpu), "--query-gpu=" + q, "--format=csv,noheader,nounits",
                                          "-lms", "100"], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.thread = threading.Thread(target=self._read, daemon=True)
            self.thread.start()
        except OSError:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([c.strip() for c in line.split(",")])

    def stop(self):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except Exception:
            self.proc.kill()
        sm, smmax, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for r in self.rows:
            try:
                sm.append(float(r[0])); smmax.append(float(r[1]))
            except (ValueError, IndexError):
                continue
            for k, name in enumerate(names):
                if len(r) > 4 + k and r[4 + k].lower().startswith("active"):
                    reasons.add(name)
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(smmax) if smmax else None,
                "reasons": sorted(reasons), "samples": len(sm)}


WORKLOADS = {
    "C2": "C2: 64-atom Si, 50 descriptors (10 G2 + 40 G4), net 50-40-40-1 tanh",
    "C3": "C3: 96-atom TiO2 (32 Ti + 64 O), 52 descriptors (8 G2 + 12 G4, two elements), two nets 52-40-40-1 tanh",
    "C4": "C4: 216-atom ternary oxide (36 Ca + 36 W + 144 O), 120 descriptors (10 G2 + 15 G4, three elements), three nets 120-60-60-60-1 tanh",
}
TOTAL_AS_NAMED = {"C2": 10000, "C3": 20000, "C4": 50000}      # BASELINE.json configs 2-4


def load_shard(ctx, capi, synth, cfg, ids, batch=1000):
    """Generate the structures with the given global ids and push them through pp_add_structures: one batched
    Set_Convert + Normalize call (pp_cart_to_frac_batch) and one descriptor launch per batch.  Returns
    (descriptor kernel ms, host seconds spent generating, host seconds in the library calls)."""
    zmap = {z: i for i, z in enumerate(cfg.elements)}
    lut = np.zeros(128, dtype=np.int32)
    for z, i in zmap.items():
        lut[z] = i
    t_desc = t_gen = t_lib = 0.0
    ids = np.asarray(ids, dtype=np.int64)
    pos = 0
    while pos < len(ids):
        chunk = ids[pos:pos + batch]
        pos += len(chunk)
        t0 = time.perf_counter()
        if len(chunk) > 1 and np.all(np.diff(chunk) == 1):
            b = synth.make_batch(cfg, int(chunk[0]), len(chunk))
        else:                                   # strided ids (s mod N sharding): one call per id, concatenated
            parts = [synth.make_batch(cfg, int(i), 1) for i in chunk]
            b = {k: np.concatenate([p_[k] for p_ in parts]) for k in parts[0]}
        t1 = time.perf_counter()
        frac = capi.cart_to_frac_batch(b["natoms"], b["cells"], b["cart"])
        ctx.add_structures(b["natoms"], b["cells"], frac, lut[b["Z"]], b["targets"], b["tags"])
        t2 = time.perf_counter()
        t_desc += ctx.phase_ms(0)
        t_gen += t1 - t0
        t_lib += t2 - t1
    return t_desc, t_gen, t_lib


def make_context(capi, synth, cfg, device):
    ctx = capi.Context(device)
    ctx.set_descriptors(len(cfg.elements), cfg.G1, cfg.G2, cfg.G3, cfg.G4, cfg.rcut)
    nets = []
    for e, z in enumerate(cfg.elements):
        net = synth.random_network(cfg, z, 1)
        ctx.set_network(e, net["sizes"], net["params"], np.zeros(cfg.NG), np.ones(cfg.NG))
        nets.append(net)
    return ctx, nets


def shard_ids(total, rank, world):
    """Setup::split_systems (Setup.cpp:582): system i goes to rank i mod Nprocs."""
    return np.arange(rank, total, world, dtype=np.int64)


FORM_NAMES = {
    2: "mlp_fused2_kernel (FP64 DMMA: forward, per-structure error, in-place deltas, weight gradients; two CTAs per SM)",
    1: "mlp_kernel<FUSED> (FP64 DMMA: forward, per-structure error, backprop, weight gradients; one CTA per SM)",
    3: "mlpw_forward_kernel + mlpw_backward_kernel (FP64 DMMA, activations kept in HBM between the phases, weights streamed "
       "into registers), one pair per element",
    0: "mlp_fused2_kernel<FORWARD> + <TRAIN> (FP64 DMMA, forward recomputed in the backward launch), one pair per element",
}


def train_bench(torch, dist, capi, synth, name, ids, rank, world, local_rank, steps, warmup, sampler=None, e2e=True, sustain_s=0.0):
    """One configuration: load this rank's structures, then time `steps` epochs device-resident (value) and through
    pp_epoch with host buffers + the host optimizer (e2e).  Returns a dict of per-rank measurements reduced over ranks."""
    cfg = synth.CONFIGS[name](len(ids))
    ctx, nets = make_context(capi, synth, cfg, local_rank)
    t0 = time.time()
    desc_ms, gen_s, lib_s = load_shard(ctx, capi, synth, cfg, ids)
    setup_wall = time.time() - t0
    if world > 1:
        if rank == 0:
            uid = torch.tensor(list(capi.Context.comm_unique_id()), dtype=torch.uint8, device="cuda")
        else:
            uid = torch.zeros(128, dtype=torch.uint8, device="cuda")
        dist.broadcast(uid, 0)
        ctx.comm_init(world, rank, bytes(uid.cpu().tolist()))
    ctx.precondition(True)
    P = ctx.n_params
    x0 = np.concatenate([n["params"] for n in nets])
    stream = torch.cuda.ExternalStream(ctx.stream)
    x_pin = torch.empty(P, dtype=torch.float64).pin_memory()
    g_pin = torch.empty(P, dtype=torch.float64).pin_memory()
    x_pin.copy_(torch.from_numpy(x0))

    def barrier():
        torch.cuda.synchronize()
        if dist is not None:
            dist.barrier()
        torch.cuda.synchronize()

    for _ in range(warmup):
        ctx.epoch_device()
    barrier()
    if sampler is not None:
        sampler.start()
    ev0 = torch.cuda.Event(enable_timing=True); ev1 = torch.cuda.Event(enable_timing=True)
    launches_before = ctx.kernel_launches
    ev0.record(stream)
    for _ in range(steps):
        ctx.epoch_device()
    ev1.record(stream)
    barrier()
    dev_ms = ev0.elapsed_time(ev1)
    launches_timed = ctx.kernel_launches - launches_before
    # per-phase device times of a few more (untimed) epochs for the roofline of the dominant kernels
    fwd_ms, bwd_ms, coll_ms = [], [], []
    for _ in range(max(3, min(steps, 10))):
        ctx.epoch_into(x_pin.data_ptr(), g_pin.data_ptr())
        fwd_ms.append(ctx.phase_ms(1)); bwd_ms.append(ctx.phase_ms(2)); coll_ms.append(ctx.phase_ms(3))
    form = ctx.epoch_form()
    comm_path = ctx.comm_path()
    e2e_ms = float("nan")
    if e2e:
        # the loop itself is the host layer's C++ training loop (what `PROPhet_b200 -train` runs): pp_epoch with HOST
        # buffers, then Optimizer::update_network + Rprop, every step
        ntrain = max(int(0.95 * len(ids) * world), 1)
        opt = _host_optimizer(x0, ntrain)
        opt.impl.train_loop(ctx.handle, x_pin.data_ptr(), g_pin.data_ptr(), warmup)
        barrier()
        e2e_s, _ = opt.impl.train_loop(ctx.handle, x_pin.data_ptr(), g_pin.data_ptr(), steps)
        torch.cuda.synchronize()
        e2e_ms = e2e_s * 1e3
    sustained = None
    if sustain_s > 0:
        # a seconds-long run of the same epochs (clocks settle under the power cap): its own clock sample
        x_pin.copy_(torch.from_numpy(x0))
        ctx.set_parameters(x0)
        n_sus = max(int(sustain_s * 1e3 / max(dev_ms / steps, 1e-3)), steps)
        if dist is not None:                    # every rank must run the SAME number of epochs (each one holds a collective)
            nt = torch.tensor([n_sus], dtype=torch.int64, device="cuda")
            dist.broadcast(nt, 0)
            n_sus = int(nt.item())
        sus_sampler = ClockSampler(local_rank) if rank == 0 else None
        barrier()
        if sus_sampler:
            sus_sampler.start()
        ev0.record(stream)
        for _ in range(n_sus):
            ctx.epoch_device()
        ev1.record(stream)
        barrier()
        sustained = {"steps": n_sus, "ms_per_step": ev0.elapsed_time(ev1) / n_sus, "clocks": sus_sampler.stop() if sus_sampler else None}
    times = torch.tensor([dev_ms, e2e_ms if e2e else 0.0, sustained["ms_per_step"] if sustained else 0.0], dtype=torch.float64, device="cuda")
    counts = torch.tensor([float(len(ids)), float(ctx.n_atoms)], dtype=torch.float64, device="cuda")
    if dist is not None:
        dist.all_reduce(times, op=dist.ReduceOp.MAX)
        dist.all_reduce(counts, op=dist.ReduceOp.SUM)
    dev_ms, e2e_ms, sus_ms = [float(v) for v in times.cpu()]
    total_struct, total_atoms = [float(v) for v in counts.cpu()]
    if sustained:
        sustained["ms_per_step"] = sus_ms
        sustained["value"] = total_struct / (sus_ms * 1e-3)
    sizes = [int(v) for v in nets[0]["sizes"]]
    F = 2 * sum(a * b for a, b in zip(sizes[:-1], sizes[1:]))
    flops_atom = 3 * F - 2 * sizes[0] * sizes[1]
    atoms_local = int(ctx.n_atoms)
    fwd, bwd = float(np.median(fwd_ms)), float(np.median(bwd_ms))
    kern_ms = bwd if form in (1, 2) else fwd + bwd               # single-pass forms: the forward phase is empty
    res = dict(cfg=cfg, name=name, P=P, form=form, ms_per_step=dev_ms / steps, value=total_struct / (dev_ms / steps * 1e-3),
               e2e_value=(total_struct / (e2e_ms / steps * 1e-3)) if e2e else None, total_struct=int(total_struct),
               total_atoms=int(total_atoms), atoms_local=atoms_local, n_local=len(ids), launches=int(launches_timed),
               flops_atom=flops_atom, kernel_ms=kern_ms, fwd_ms=fwd, bwd_ms=bwd, coll_ms=float(np.median(coll_ms)),
               desc_ms=desc_ms, gen_s=gen_s, lib_s=lib_s, setup_wall=setup_wall, sustained=sustained, sizes=sizes, comm_path=comm_path)
    ctx.close()
    return res


def roofline_of(res, peaks):
    achieved = res["flops_atom"] * res["atoms_local"] / (res["kernel_ms"] * 1e-3) / 1e12
    cfg = res["cfg"]
    return {"bound": "tensor", "kernel": FORM_NAMES.get(res["form"], "?"), "achieved": achieved, "peak": peaks["fp64_dmma_tflops"],
            "unit": "TFLOP/s", "frac": achieved / peaks["fp64_dmma_tflops"], "peak_source": peaks["source"],
            "algorithmic_flops_per_atom": res["flops_atom"], "atoms_per_launch": res["atoms_local"], "kernel_ms": res["kernel_ms"],
            "traffic": _ncu_traffic(res["name"], res["n_local"], res["form"]),
            "hbm_view": {"algorithmic_bytes_per_atom": 8 * cfg.NG,
                         "achieved_gbs": 8 * cfg.NG * res["atoms_local"] / (res["kernel_ms"] * 1e-3) / 1e9, "peak_gbs": _measured_hbm()}}


def multirank_parity(torch, dist, capi, synth, rank, world, local_rank, name="C2", n_union=64):
    """N > 1 only, outside every timed region: a small union of structures sharded s mod N (Setup.cpp:582) must give, after
    the all-reduce, the [gradient, SSE, vSSE, nval] of a single-rank epoch over the same structures."""
    cfg = synth.CONFIGS[name](n_union)
    ctx, nets = make_context(capi, synth, cfg, local_rank)
    load_shard(ctx, capi, synth, cfg, shard_ids(n_union, rank, world))
    uid = torch.tensor(list(capi.Context.comm_unique_id()) if rank == 0 else [0] * 128, dtype=torch.uint8, device="cuda")
    dist.broadcast(uid, 0)
    ctx.comm_init(world, rank, bytes(uid.cpu().tolist()))
    x0 = np.concatenate([n["params"] for n in nets])
    r = ctx.epoch(x0)
    ctx.close()
    out = None
    if rank == 0:
        one, _ = make_context(capi, synth, cfg, local_rank)
        load_shard(one, capi, synth, cfg, np.arange(n_union))
        r1 = one.epoch(x0)
        one.close()
        gmax = float(np.max(np.abs(r1["gradient"])))
        out = {"workload": "%s, %d structures sharded s mod %d vs one rank" % (name, n_union, world),
               "gradient_rel_err": float(np.max(np.abs(r["gradient"] - r1["gradient"])) / gmax),
               "sse_rel_err": abs(r["SSE"] - r1["SSE"]) / abs(r1["SSE"]), "vsse_rel_err": abs(r["vSSE"] - r1["vSSE"]) / max(abs(r1["vSSE"]), 1e-300),
               "nval_equal": bool(r["nval"] == r1["nval"]), "tolerance": 1e-12}
        out["ok"] = bool(out["gradient_rel_err"] <= 1e-12 and out["sse_rel_err"] <= 1e-12 and out["vsse_rel_err"] <= 1e-12 and out["nval_equal"])
    return out


# ------------------------------------------------------------------------------------------------------------
def run_ours(args):
    import torch
    from prophet_b200 import capi, synth

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if world != args.gpus:
        if rank == 0 and world != 1:
            print("warning: WORLD_SIZE=%d but --gpus %d" % (world, args.gpus), file=sys.stderr)
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device (this benchmark has no CPU fallback; use --impl reference for the CPU arm)")
    torch.cuda.set_device(local_rank)
    dist = None
    if world > 1:
        import torch.distributed as dist_
        dist = dist_
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
        os.environ.setdefault("PP_NCCL_LIB", _torch_nccl_path())

    name = args.config
    # headline: weak scaling, `--structures` per GPU (C2: 10 000 = the configuration as named on one GPU)
    n_local = args.structures if args.structures > 0 else TOTAL_AS_NAMED[name]
    sampler = ClockSampler(local_rank) if rank == 0 else None
    head = train_bench(torch, dist, capi, synth, name, np.arange(rank * n_local, (rank + 1) * n_local), rank, world, local_rank,
                       args.steps, args.warmup, sampler=sampler, e2e=True, sustain_s=args.sustain)
    clocks = sampler.stop() if rank == 0 else None
    peaks = _fp64_peaks()
    out = None
    if rank == 0:
        cfg = head["cfg"]
        out = {
            "metric": METRIC, "value": head["value"], "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": head["ms_per_step"], "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "f64", "data": "synthetic",
            "config": {"workload": "%s, %d structures per GPU, one epoch" % (WORKLOADS[name], n_local),
                       "structures_total": head["total_struct"], "parallelism": "dp%d" % world,
                       "l2": "inputs (%.0f MB FP64 descriptors per GPU) larger than L2, no flush" % (head["atoms_local"] * cfg.NG * 8 / 1e6)},
            "e2e": {"value": head["e2e_value"], "unit": UNIT, "h2d_bytes_per_step": 8 * head["P"], "d2h_bytes_per_step": 8 * (head["P"] + 3),
                    "includes": "the host layer's C++ training loop: pp_epoch with pinned host buffers + Optimizer::update_network (Rprop) per step"},
            "gpu_launches": head["launches"],
            "clocks": clocks,
            "roofline": roofline_of(head, peaks),
            "phases_ms": {"forward": head["fwd_ms"], "backward": head["bwd_ms"], "reduce_and_collective": head["coll_ms"]},
            "collective_path": head["comm_path"],
            "setup": {"descriptor_kernel_ms_total": head["desc_ms"], "structures_per_s_kernel": n_local / (head["desc_ms"] * 1e-3),
                      "wall_s": head["setup_wall"], "host_generation_s": head["gen_s"], "library_calls_s": head["lib_s"],
                      "structures_per_s_load": n_local / max(head["lib_s"], 1e-9),
                      "note": "library_calls_s = pp_cart_to_frac_batch + pp_add_structures (staging, upload, descriptor kernel) per batch of 1000"},
        }
        if head["sustained"]:
            out["sustained"] = head["sustained"]
        if world == 1 and not args.no_cpu_baseline:
            out["cpu_baseline"] = reference_epoch_rate(cores=1, structures_per_core=args.cpu_sample, iters=6, config=name)
    # the other training configurations BASELINE.json names, each AS NAMED: its total sharded s mod N over the GPUs
    # (strong scaling: at N = 1 the whole set is on one GPU)
    if not args.no_other_configs:
        others = {}
        for oname in ("C2", "C3", "C4"):
            if oname == name and world == 1 and n_local == TOTAL_AS_NAMED[oname]:
                continue                         # identical to the headline run
            total = int(TOTAL_AS_NAMED[oname] * args.other_scale)
            r = train_bench(torch, dist, capi, synth, oname, shard_ids(total, rank, world), rank, world, local_rank,
                            max(args.steps // 2, 5), args.warmup, e2e=True)
            if rank == 0:
                others[oname] = {"workload": "%s, %d structures in total sharded s mod %d (Setup.cpp:582), one epoch" % (WORKLOADS[oname], total, world),
                                 "scaling": "strong", "structures_total": r["total_struct"], "ms_per_step": r["ms_per_step"], "value": r["value"],
                                 "unit": UNIT, "e2e": {"value": r["e2e_value"], "unit": UNIT, "h2d_bytes_per_step": 8 * r["P"],
                                                       "d2h_bytes_per_step": 8 * (r["P"] + 3)},
                                 "epoch_form": r["form"], "parameters": r["P"], "gpu_launches": r["launches"],
                                 "phases_ms": {"forward": r["fwd_ms"], "backward": r["bwd_ms"], "reduce_and_collective": r["coll_ms"]},
                                 "roofline": roofline_of(r, peaks),
                                 "setup": {"descriptor_kernel_ms_total": r["desc_ms"], "host_generation_s": r["gen_s"], "library_calls_s": r["lib_s"],
                                           "structures_per_s_load": r["n_local"] / max(r["lib_s"], 1e-9)}}
        if rank == 0:
            out["other_configs"] = others
    if world > 1:
        par = multirank_parity(torch, dist, capi, synth, rank, world, local_rank)
        if rank == 0:
            out["multi_rank_parity"] = par
            assert par["ok"], "multi-rank epoch differs from the single-rank epoch: %r" % (par,)
    if not args.no_md:
        # pair_nn (second metric of BASELINE.json); at N > 1 every rank runs an independent replica of the box
        md = run_md_bench(torch, capi, synth, local_rank, args.md_cells, args.md_steps, 2, dist=dist, rank=rank, world=world)
        if dist is not None:
            t = torch.tensor([md["ms_per_step"], md["e2e"]["ms_per_step"]], dtype=torch.float64, device="cuda")
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            md["ms_per_step"], md["e2e"]["ms_per_step"] = [float(v) for v in t.cpu()]
            md["value"] = world * md["atoms_local"] / (md["ms_per_step"] * 1e-3)
            md["e2e"]["value"] = world * md["atoms_local"] / (md["e2e"]["ms_per_step"] * 1e-3)
            md["n_gpus"] = world
            md["partition"] = "%d sub-boxes along x (weak scaling), ghost forces of the x faces exchanged with ring neighbours " \
                              "(NCCL send/recv), y/z images folded locally, energy all-reduced; e2e leg: per-rank host round trip " \
                              "without the exchange" % world
        if rank == 0:
            if world == 1 and not args.no_cpu_baseline:
                md["cpu_baseline"] = reference_md_rate()
            out["pair_nn"] = md
    if rank == 0:
        emit(out)
    if dist is not None:
        dist.destroy_process_group()


# ------------------------------------------------------------------------------------------------------------
# second headline metric: pair_nn atom-steps/s (C5: diamond-Si supercell, C2 rows, 50-40-40-1 net)
def md_system(synth, cells, seed=5, skin=1.0, with_owner=False):
    cfg = synth.config_c5(cells)
    cell, cart, Z, _, _ = synth.make_structure(cfg, seed)
    if with_owner:
        x, Zall, nlocal, owner, shift = synth.expand_ghosts(cell, cart, Z, cfg.rcut + skin, with_owner=True)
        return cfg, np.ascontiguousarray(x), nlocal, owner, shift
    x, Zall, nlocal = synth.expand_ghosts(cell, cart, Z, cfg.rcut + skin)
    return cfg, np.ascontiguousarray(x), nlocal


def run_md_bench(torch, capi, synth, device, cells, steps, warmup, dist=None, rank=0, world=1):
    """pp_md_build_neighbors once (LAMMPS rebuilds every 10 steps, pair_nn.cpp:276), then `steps` calls of the
    PairNN::compute equivalent: device-resident (value) and through host buffers (e2e).

    N > 1 (weak scaling): the global box is N sub-boxes side by side along x, one per GPU, periodic.  Each rank
    computes forces on its locals AND ghosts exactly as the reference does (pair_nn.cpp:182 writes ghost forces and
    relies on LAMMPS' reverse communication); the harness then plays LAMMPS' part: ghost forces of y/z images are
    folded onto their owners locally, those of the two x faces travel to the neighbouring ranks (NCCL send/recv,
    ring) and are added there, and the energy is all-reduced.  That exchange is inside the timed region."""
    cfg, x, nlocal, owner, shift = md_system(synth, cells, with_owner=True)
    ntotal = len(x)
    ctx = capi.Context(device)
    ctx.set_descriptors(1, cfg.G1, cfg.G2, cfg.G3, cfg.G4, cfg.rcut)
    net = synth.random_network(cfg, 14, 1)
    mean = 0.1 * synth.normal(9, cfg.NG, 3); scale = 0.5 + np.abs(synth.normal(9, cfg.NG, 4))
    ctx.set_network(0, net["sizes"], net["params"], mean, scale)
    ctx.md_set_types([0, 0])
    types = np.ones(ntotal, dtype=np.int32)
    t0 = time.perf_counter()
    ctx.md_build_neighbors(x, types, nlocal, cfg.rcut + 1.0)
    build_s = time.perf_counter() - t0
    stream = torch.cuda.ExternalStream(ctx.stream)
    dev = torch.device("cuda", device)
    xd = torch.from_numpy(x).to(dev)
    fd = torch.zeros_like(xd)
    ed = torch.zeros(1, dtype=torch.float64, device=dev)
    torch.cuda.synchronize()
    exchange = None
    if world > 1:
        own = torch.from_numpy(owner).to(dev)
        gidx = torch.arange(nlocal, ntotal, device=dev)
        sx = torch.from_numpy(shift[:, 0].copy()).to(dev)
        sel_self, sel_left, sel_right = (sx == 0), (sx < 0), (sx > 0)
        g_self, o_self = gidx[sel_self], own[sel_self]
        g_left, o_left = gidx[sel_left], own[sel_left]
        g_right, o_right = gidx[sel_right], own[sel_right]
        left, right = (rank - 1) % world, (rank + 1) % world
        recv_r = torch.empty((len(g_left), 3), dtype=torch.float64, device=dev)    # from the right neighbour: its left ghosts = my atoms
        recv_l = torch.empty((len(g_right), 3), dtype=torch.float64, device=dev)

        def exchange():
            with torch.cuda.stream(stream):
                fd.index_add_(0, o_self, fd[g_self])                      # local fold (y/z periodic images)
                send_l, send_r = fd[g_left].contiguous(), fd[g_right].contiguous()
                ops = [dist.P2POp(dist.isend, send_l, left), dist.P2POp(dist.isend, send_r, right),
                       dist.P2POp(dist.irecv, recv_r, right), dist.P2POp(dist.irecv, recv_l, left)]
                for w_ in dist.batch_isend_irecv(ops):
                    w_.wait()
                fd.index_add_(0, o_left, recv_r)
                fd.index_add_(0, o_right, recv_l)
                dist.all_reduce(ed)

    def one_step():
        with torch.cuda.stream(stream):
            fd.zero_()                                                    # LAMMPS clears forces before Pair::compute
        ctx.md_step_device(xd.data_ptr(), fd.data_ptr(), ed.data_ptr())
        if exchange is not None:
            exchange()

    for _ in range(warmup):
        one_step()
    torch.cuda.synchronize()
    if dist is not None:
        dist.barrier()
        torch.cuda.synchronize()
    ev0 = torch.cuda.Event(enable_timing=True); ev1 = torch.cuda.Event(enable_timing=True)
    l0 = ctx.kernel_launches
    ev0.record(stream)
    for _ in range(steps):
        one_step()
    ev1.record(stream)
    ev1.synchronize()
    torch.cuda.synchronize()
    dev_ms = ev0.elapsed_time(ev1) / steps
    launches = ctx.kernel_launches - l0
    # property check (outside the timed region): once every ghost force is back on its owner the forces on the local
    # atoms are those of the periodic system, whose total force vanishes
    one_step()
    with torch.cuda.stream(stream):
        f_loc = fd[:nlocal].clone()
        if world == 1:
            f_loc.index_add_(0, torch.from_numpy(owner).to(dev), fd[nlocal:])
        fsum = float(f_loc.sum(dim=0).abs().max() / f_loc.abs().max())
    torch.cuda.synchronize()
    # e2e: pinned host positions in, forces / energies out, every step
    x_pin = torch.from_numpy(x).pin_memory()
    f_pin = torch.zeros(ntotal, 3, dtype=torch.float64).pin_memory()
    ea_pin = torch.zeros(max(nlocal, 1), dtype=torch.float64).pin_memory()
    for _ in range(max(1, warmup - 1)):
        ctx.md_step_ptr(x_pin.data_ptr(), f_pin.data_ptr(), ea_pin.data_ptr())
    phases = []
    t0 = time.perf_counter()
    for _ in range(steps):
        ctx.md_step_ptr(x_pin.data_ptr(), f_pin.data_ptr(), ea_pin.data_ptr())
        phases.append((ctx.phase_ms(4), ctx.phase_ms(5), ctx.phase_ms(6)))
    e2e_ms = (time.perf_counter() - t0) * 1e3 / steps
    ph = np.median(np.array(phases), axis=0)
    # algorithmic flops per atom-step (measured mean neighbour count n)
    off, _nb = None, None
    nneigh = None
    out = {"metric": "pair_nn atom-steps/sec", "value": nlocal / (dev_ms * 1e-3), "unit": "atom-steps/s", "ms_per_step": dev_ms,
           "atoms_local": int(nlocal), "atoms_ghost": int(ntotal - nlocal), "steps": steps, "warmup": warmup,
           "workload": "C5: diamond Si %dx%dx%d cells, Rcut 6 + skin 1, 50 descriptors, net 50-40-40-1" % (cells, cells, cells),
           "e2e": {"value": nlocal / (e2e_ms * 1e-3), "unit": "atom-steps/s", "ms_per_step": e2e_ms,
                   "h2d_bytes_per_step": 2 * 24 * ntotal, "d2h_bytes_per_step": 24 * ntotal + 8 * nlocal + 8},
           "phases_ms": {"descriptors": float(ph[0]), "network": float(ph[1]), "forces": float(ph[2])},
           "neighbor_build_s": build_s, "gpu_launches": int(launches), "total_force_over_max_force": fsum}
    peaks = _fp64_peaks()
    flops_desc, flops_force = 0.41e6, 1.21e6      # per atom at n = 45; the dominant kernel is the force kernel
    out["roofline"] = {"bound": "fp64-fma", "kernel": "md_force_kernel", "achieved": flops_force * nlocal / (ph[2] * 1e-3) / 1e12,
                       "peak": peaks["fp64_fma_tflops"], "unit": "TFLOP/s",
                       "frac": flops_force * nlocal / (ph[2] * 1e-3) / 1e12 / peaks["fp64_fma_tflops"],
                       "algorithmic_flops_per_atom": flops_force, "peak_source": peaks["source"],
                       # SURVEY 8(d) counts every G4 row per pair (12 + 30 x 40 flops); the factorised power-basis evaluation
                       # EXECUTES far fewer, so the fraction above can exceed 1.  What the kernel really issues (ncu,
                       # profiles/r01_ncu_md_final.txt capture: dfma x 2 + dmul + dadd = 0.284 MFLOP per atom at C5's rows):
                       "executed": {"flops_per_atom": 0.284e6, "achieved": 0.284e6 * nlocal / (ph[2] * 1e-3) / 1e12,
                                    "frac": 0.284e6 * nlocal / (ph[2] * 1e-3) / 1e12 / peaks["fp64_fma_tflops"],
                                    "source": "ncu sm__sass_thread_inst_executed_op_{dfma,dmul,dadd}_pred_on of md_force_kernel"}}
    ctx.close()
    return out


def reference_md_rate(cells=2, cores=1):
    """atom-steps/s of the reference's PairNN::compute call chain (oracle/_ref/ref_driver md): `cores` independent
    processes, each one step on its own copy of a small box (what LAMMPS' spatial decomposition gives every MPI rank)."""
    from prophet_b200 import ckpt, synth
    exe = os.path.join(ROOT, "oracle", "_ref", "ref_driver")
    if not os.path.exists(exe):
        return {"value": None, "unit": "atom-steps/s", "cores": cores, "kind": "reference", "sample": "oracle/_ref/ref_driver missing"}
    cfg, x, nlocal = md_system(synth, cells)
    numneigh, offsets, flat = synth.brute_full_neighbor_list(x, nlocal, cfg.rcut + 1.0)
    net = synth.random_network(cfg, 14, 1)
    d = tempfile.mkdtemp(prefix="ppmd_")
    secs = []
    try:
        ckpt.write_checkpoint(os.path.join(d, "pot_Si"), "Si", ["Si"], {"G1": [], "G2": cfg.G2, "G3": [], "G4": cfg.G4},
                              np.zeros(cfg.NG), np.ones(cfg.NG), 0.0, 1.0, net["sizes"], net["params"], header_precision=17)
        with open(os.path.join(d, "case.txt"), "w") as f:
            f.write("Rcut %r\nntypes 1\ncoeff 1 pot_Si\natoms %d %d\n" % (float(cfg.rcut), nlocal, len(x) - nlocal))
            for r in x:
                f.write("1 %r %r %r\n" % (float(r[0]), float(r[1]), float(r[2])))
            for i in range(nlocal):
                f.write("neigh %d %d %s\n" % (i, numneigh[i], " ".join(str(j) for j in flat[offsets[i]:offsets[i + 1]])))
        procs = [subprocess.Popen([exe, "md", "case.txt", "out%d.bin" % c], cwd=d, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
                 for c in range(cores)]
        for pr in procs:
            so, _ = pr.communicate()
            m = re.search(r"MD_COMPUTE_SECONDS ([0-9.eE+-]+)", so)
            if m:
                secs.append(float(m.group(1)))
    finally:
        shutil.rmtree(d, ignore_errors=True)
    value = sum(nlocal / t for t in secs) if secs else None
    return {"metric": "pair_nn atom-steps/sec", "value": value, "unit": "atom-steps/s", "cores": cores, "kind": "reference",
            "sample": "%d process(es) x 1 step, %d local + %d ghost Si atoms each, unmodified reference classes driven as PairNN::compute "
                      "does (oracle/_ref/ref_driver md)" % (len(secs), nlocal, len(x) - nlocal),
            "seconds_per_process": float(np.mean(secs)) if secs else None}


class _host_optimizer:
    """Host-side optimizer step of the e2e loop: the C++ host layer's Optimizer::update_network + Rprop
    (prophet_b200/host/pph_optimizer.cpp), reading the gradient from and writing the new parameters to pinned memory."""

    def __init__(self, x0, ntrain):
        from prophet_b200 import hostapi
        self.impl = hostapi.Optimizer(x0, "rprop", max(ntrain, 1))

    def step(self, x_pin, g_pin, sse=0.0):
        self.impl.step_ptr(x_pin.data_ptr(), g_pin.data_ptr(), sse)


def _torch_nccl_path():
    import glob
    import torch
    base = os.path.dirname(os.path.dirname(torch.__file__))
    for pat in ("nvidia/nccl/lib/libnccl.so.2", "torch/lib/libnccl.so.2"):
        hits = glob.glob(os.path.join(base, pat))
        if hits:
            return hits[0]
    return "libnccl.so.2"


def _ncu_traffic(name, n_struct, form=1):
    """dram__bytes_read.sum + dram__bytes_write.sum of the dominant kernel(s), per epoch, from the committed ncu captures
    (only valid for the workload they were taken on; null otherwise)."""
    key = {2: "mlp_fused2_kernel_%s_%d", 1: "mlp_kernel_fused_%s_%d", 3: "mlpw_kernels_%s_%d", 0: "two_phase_recompute_%s_%d"}[form] % (name, n_struct)
    for fn in ("r02_traffic.json", "r01_traffic.json"):
        try:
            with open(os.path.join(ROOT, "profiles", fn)) as f:
                d = json.load(f).get(key)
            if d:
                return d["dram_bytes_read"] + d["dram_bytes_write"]
        except Exception:
            pass
    return None


def _measured_hbm():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            return json.load(f)["hbm_gbs"]
    except Exception:
        return 6650.0


def _fp64_peaks():
    """FP64 peaks are not in MEASURED_PEAKS.json (HBM and bf16 only); tools/peaks_microbench.cu measured them on this
    pool's B200 (profiles/r01_peaks_microbench.json)."""
    try:
        with open(os.path.join(ROOT, "profiles", "r01_peaks_microbench.json")) as f:
            d = json.load(f)
        return {"fp64_dmma_tflops": d["fp64_dmma_m8n8k4_tflops"], "fp64_fma_tflops": d["fp64_fma_tflops"],
                "source": "measured: tools/peaks_microbench.cu on B200 (profiles/r01_peaks_microbench.json)"}
    except Exception:
        return {"fp64_dmma_tflops": 37.0, "fp64_fma_tflops": 33.5, "source": "fallback constants"}


# ------------------------------------------------------------------------------------------------------------
# The reference's own CPU implementation (oracle/_ref/PROPhet = unmodified reference CLI, built by oracle/build_ref.sh)
def _write_ref_case(cwd, synth, first, count, iters, config="C2"):
    cfg = synth.CONFIGS[config](count)
    structs = [synth.make_structure(cfg, first + s) for s in range(count)]
    synth.write_prophet_xml(os.path.join(cwd, "PROPhet.xml"), structs)
    synth.write_input_file(os.path.join(cwd, "input_file"), cfg, niterations=iters, extra=("seed = 1",))


def _run_ref_timed(cwd, iters):
    """Run the reference CLI with line-buffered stdout and time the iteration lines (Nprint = 1)."""
    exe = os.path.join(ROOT, "oracle", "_ref", "PROPhet")
    cmd = ["stdbuf", "-oL", exe, "-in", "input_file", "-train"] if shutil.which("stdbuf") else [exe, "-in", "input_file", "-train"]
    p = subprocess.Popen(cmd, cwd=cwd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
    stamps = []
    it_line = re.compile(r"^\d+\s+[-+0-9.eE]+\s+[-+0-9.eE]+(\s+[-+0-9.eE]+)?\s*$")   # Optimizer.cpp:980-988
    for line in p.stdout:
        if it_line.match(line):
            stamps.append(time.perf_counter())
    p.wait()
    return stamps


def reference_epoch_rate(cores, structures_per_core, iters, steps=None, warmup=0, config="C2"):
    """structures/s/epoch of the reference: `cores` independent serial processes on disjoint shards (the reference's
    own round-robin split, Setup.cpp:582, without the microsecond-scale Reduce/Bcast; no MPI in this image)."""
    from prophet_b200 import synth
    exe = os.path.join(ROOT, "oracle", "_ref", "PROPhet")
    if not os.path.exists(exe):
        return {"value": None, "unit": UNIT, "cores": cores, "kind": "reference", "sample": "oracle/_ref/PROPhet missing"}
    dirs = []
    results = [None] * cores
    try:
        for c in range(cores):
            d = tempfile.mkdtemp(prefix="ppref_%d_" % c)
            dirs.append(d)
            _write_ref_case(d, synth, c * structures_per_core, structures_per_core, iters, config)

        def work(c):
            results[c] = _run_ref_timed(dirs[c], iters)

        threads = [threading.Thread(target=work, args=(c,)) for c in range(cores)]
        for t in threads:
            t.start()
        for t in threads:
            t.join()
    finally:
        for d in dirs:
            shutil.rmtree(d, ignore_errors=True)
    # per-process epoch time from the gaps between iteration lines (skip `warmup` gaps)
    rates, epoch_ms = [], []
    for st in results:
        if not st or len(st) < 2 + warmup:
            continue
        gaps = np.diff(st)[warmup:]
        if steps:
            gaps = gaps[:steps]
        dt = float(np.mean(gaps))
        epoch_ms.append(dt * 1e3)
        rates.append(structures_per_core / dt)
    value = float(np.sum(rates)) if rates else None
    return {"value": value, "unit": UNIT, "cores": cores, "kind": "reference",
            "sample": "%d processes x %d %s structures, %d epochs each, unmodified reference CLI (oracle/_ref/PROPhet -train), "
                      "epoch time from iteration-line timestamps" % (cores, structures_per_core, config, iters),
            "epoch_ms_per_process": float(np.mean(epoch_ms)) if epoch_ms else None}


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    cores = os.cpu_count() or 1
    # bounded sample per host core, sized so that the run ends within minutes whatever the configuration
    per_core = args.cpu_sample if args.config == "C2" else max(2, args.cpu_sample // (4 if args.config == "C3" else 12))
    iters = args.warmup + args.steps + 1
    r = reference_epoch_rate(cores, per_core, iters, steps=args.steps, warmup=args.warmup, config=args.config)
    ms = r.get("epoch_ms_per_process")
    out = {"impl": "reference", "metric": METRIC, "value": r["value"], "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
           "warmup": args.warmup, "ms_per_step": ms, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
           "dtype": "f64", "data": "synthetic",
           "config": {"workload": "%s, one epoch; bounded sample of %d structures per host core" % (WORKLOADS[args.config], per_core),
                      "parallelism": "cpu x%d" % cores},
           "cpu_baseline": r,
           "e2e": {"value": r["value"], "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
           "gpu_launches": 0}
    if not args.no_md:
        # second metric: the reference's PairNN::compute call chain on every host core (disjoint copies of a small box)
        out["pair_nn"] = reference_md_rate(cores=cores)
    emit(out)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--config", default="C2", choices=["C2", "C3", "C4"], help="training configuration of the headline line")
    ap.add_argument("--structures", type=int, default=0, help="structures per GPU of the headline run (0 = the configuration's total as named)")
    ap.add_argument("--no-other-configs", action="store_true", help="skip the other_configs object (C2 / C3 / C4 as named, sharded over the GPUs)")
    ap.add_argument("--other-scale", type=float, default=1.0, help="fraction of the named totals used for other_configs (testing)")
    ap.add_argument("--sustain", type=float, default=2.0, help="seconds of back-to-back epochs for the sustained figure (0 = skip)")
    ap.add_argument("--cpu-sample", type=int, default=24, help="structures per host core for the CPU arm")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-md", action="store_true", help="skip the pair_nn (MD) section")
    ap.add_argument("--md-cells", type=int, default=50, help="diamond-Si supercell edge in unit cells (C5 = 50 -> 1M atoms)")
    ap.add_argument("--md-steps", type=int, default=5)
    args = ap.parse_args()
    if args.impl == "reference":
        run_reference(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
