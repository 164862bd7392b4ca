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


def make_context(capi, synth, cfg, device, fp_mode=64):
    ctx = capi.Context(device, fp_mode)
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
    5: "mlp_fused2_kernel<FORWARD> + <TRAIN> (FP64 DMMA, two CTAs per SM; activations and their derivatives kept in HBM between the "
       "two launches), one pair per element",
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
    if world > 1:
        # the ranks leave the host barrier up to hundreds of microseconds apart; one more untimed epoch (whose sum over
        # the ranks is a device-side rendezvous) lines the GPUs' timelines up, so that the K timed epochs that follow do
        # not include the first rank's wait for the last one to start
        ctx.epoch_device()
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
        if world > 1:
            opt.impl.train_loop(ctx.handle, x_pin.data_ptr(), g_pin.data_ptr(), 1)      # lines the ranks up (see above)
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
        if world > 1:
            ctx.epoch_device()
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
                         # stored two-phase forms: H_l and T'_l of every hidden node written once and read once
                         "stored_activation_bytes_per_atom": 32 * sum(res["sizes"][1:-1]) if res["form"] in (3, 5) else 0,
                         "achieved_gbs": 8 * cfg.NG * res["atoms_local"] / (res["kernel_ms"] * 1e-3) / 1e9, "peak_gbs": _measured_hbm()}}


def fp32_mode_bench(torch, capi, synth, device, steps, warmup, n_struct=10000, n_check=512):
    """north_star's FP32 mode (pp_create(..., fp_mode = 32), tolerance relative 1e-5 against the FP64 path) on the headline
    workload, one GPU: the epoch time of the FP32 kernel (FP32 descriptors in HBM, forward / delta products as 3xTF32 on
    tcgen05.mma with TMEM accumulators, FP64 energies and sums) and, outside the timed region, its gradient and SSE
    against the FP64 kernels on the same structures and parameters."""
    cfg = synth.CONFIGS["C2"](n_struct)
    ctx, nets = make_context(capi, synth, cfg, device, 32)
    load_shard(ctx, capi, synth, cfg, np.arange(n_struct))
    ctx.precondition(True)
    x0 = np.concatenate([n["params"] for n in nets])
    ctx.set_parameters(x0)
    stream = torch.cuda.ExternalStream(ctx.stream)
    for _ in range(max(warmup, 3)):
        ctx.epoch_device()
    torch.cuda.synchronize()
    ev0 = torch.cuda.Event(enable_timing=True); ev1 = torch.cuda.Event(enable_timing=True)
    ev0.record(stream)
    for _ in range(steps):
        ctx.epoch_device()
    ev1.record(stream)
    torch.cuda.synchronize()
    ms = ev0.elapsed_time(ev1) / steps
    form = ctx.epoch_form()
    atoms = int(ctx.n_atoms)
    ctx.close()
    # parity on a subset: same structures, same parameters, the two modes
    cfgc = synth.CONFIGS["C2"](n_check)
    res = {}
    for mode in (64, 32):
        c, _ = make_context(capi, synth, cfgc, device, mode)
        load_shard(c, capi, synth, cfgc, np.arange(n_check))
        c.precondition(True)
        res[mode] = c.epoch(x0)
        c.close()
    gmax = float(np.max(np.abs(res[64]["gradient"])))
    grad_err = float(np.max(np.abs(res[32]["gradient"] - res[64]["gradient"])) / gmax)
    sse_err = abs(res[32]["SSE"] - res[64]["SSE"]) / abs(res[64]["SSE"])
    sizes = [int(v) for v in nets[0]["sizes"]]
    F = 2 * sum(a * b for a, b in zip(sizes[:-1], sizes[1:]))
    flops_atom = 3 * F - 2 * sizes[0] * sizes[1]
    bytes_atom = 4 * ((cfg.NG + 3) // 4 * 4)
    return {"workload": "%s, %d structures, fp_mode 32" % (WORKLOADS["C2"], n_struct), "dtype": "f32 (3xTF32 products, FP32 accumulation; FP64 energies, errors, sums)",
            "ms_per_step": ms, "value": n_struct / (ms * 1e-3), "unit": "structures/s", "steps": steps, "epoch_form": form,
            "kernel": "mlp32_umma_kernel (tcgen05.mma kind::tf32 from shared memory into TMEM for the forward and delta products, "
                      "tcgen05.ld epilogues; weight gradients on mma.sync tf32)",
            "parity_vs_fp64": {"structures": n_check, "gradient_rel_err": grad_err, "sse_rel_err": sse_err, "tolerance": 1e-5,
                               "ok": bool(grad_err <= 1e-5 and sse_err <= 1e-5)},
            "roofline": {"bound": "hbm", "achieved": bytes_atom * atoms / (ms * 1e-3) / 1e9, "peak": _measured_hbm(), "unit": "GB/s",
                         "frac": bytes_atom * atoms / (ms * 1e-3) / 1e9 / _measured_hbm(), "algorithmic_bytes_per_atom": bytes_atom,
                         "note": "FP32 descriptors are half the bytes; the epoch is far from the HBM bound -- it is bound by the legacy mma.sync "
                                 "tf32 rate of the weight-gradient products and by the serial phases of one resident CTA per SM",
                         "algorithmic_tflops": flops_atom * atoms / (ms * 1e-3) / 1e12}}


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
                       "l2": "inputs (%.0f MB FP64 descriptors per GPU) larger than L2, no flush" % (head["atoms_local"] * cfg.NG * 8 / 1e6),
                       "timing": "CUDA events on the library's stream round exactly K epochs, max over ranks"
                                 + ("; at N > 1 one untimed epoch follows the host barrier so that the ranks start the timed epochs together"
                                    if world > 1 else "")},
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
    if world == 1 and not args.no_fp32:
        # an extra object, not the headline: a failure here is reported in the line instead of costing the line
        try:
            out["fp32"] = fp32_mode_bench(torch, capi, synth, local_rank, args.steps, args.warmup)
        except Exception as exc:      # noqa: BLE001
            out["fp32"] = {"error": "%s: %s" % (type(exc).__name__, exc)}
    if world > 1:
        par = multirank_parity(torch, dist, capi, synth, rank, world, local_rank)
        if rank == 0:
            out["multi_rank_parity"] = par
            assert par["ok"], "multi-rank epoch differs from the single-rank epoch: %r" % (par,)
    if not args.no_md:
        # pair_nn (second metric of BASELINE.json): one box cut over the GPUs, exchange owned by the library
        md = run_md_bench(torch, capi, synth, local_rank, args.md_cells, args.md_steps, 2, dist=dist, rank=rank, world=world)
        if world > 1:
            t = torch.tensor(list(capi.Context.comm_unique_id()) if rank == 0 else [0] * 128, dtype=torch.uint8, device="cuda")
            dist.broadcast(t, 0)
            chk = md_partition_check(torch, capi, synth, dist, local_rank, rank, world, bytes(t.cpu().tolist()))
            if rank == 0:
                md["partition_parity"] = chk
                assert chk["ok"], "partitioned forces differ from the single-rank forces: %r" % (chk,)
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


def md_partition(torch, capi, synth, dist, device, cells, rank, world, comm_uid=None, seed=5, skin=1.0):
    """This rank's slab of the global diamond-Si box (cells^3 unit cells, periodic), set up in a context: atoms uploaded,
    full neighbour list built on the GPU, ghost-force exchange lists installed.  Returns (ctx, cfg, partition dict)."""
    cfg = synth.config_c5(cells)
    cell, cart, Z, _, _ = synth.make_structure(cfg, seed)          # the same global box on every rank (seeded)
    part = synth.slab_partition(cell, cart, rank, world, cfg.rcut + skin)
    x, nlocal = part["x"], part["nlocal"]
    ntotal = len(x)
    ctx = capi.Context(device)
    ctx.set_descriptors(1, cfg.G1, cfg.G2, cfg.G3, cfg.G4, cfg.rcut)
    net = synth.random_network(cfg, 14, 1)
    mean = 0.1 * synth.normal(9, cfg.NG, 3); scale = 0.5 + np.abs(synth.normal(9, cfg.NG, 4))
    ctx.set_network(0, net["sizes"], net["params"], mean, scale)
    ctx.md_set_types([0, 0])
    t0 = time.perf_counter()
    ctx.md_build_neighbors(x, np.ones(ntotal, dtype=np.int32), nlocal, cfg.rcut + skin)
    part["build_s"] = time.perf_counter() - t0
    if world > 1:
        ctx.comm_init(world, rank, comm_uid)
    # exchange lists: ghosts owned by this rank fold locally; the others go to their owners, who need to know which of
    # THEIR atoms each value belongs to (set-up time exchange of index lists, as LAMMPS' Comm::borders does)
    gidx = np.arange(nlocal, ntotal, dtype=np.int32)
    mine = part["ghost_rank"] == rank
    wanted = {}
    for o in np.unique(part["ghost_rank"]):
        if int(o) != rank:
            sel = part["ghost_rank"] == o
            wanted[int(o)] = (gidx[sel], part["ghost_local"][sel])
    peers = []
    if world > 1:
        gathered = [None] * world
        dist.all_gather_object(gathered, {o: v[1] for o, v in wanted.items()})
        for o in sorted(set(wanted) | {r for r in range(world) if r != rank and rank in gathered[r]}):
            send_idx = wanted[o][0] if o in wanted else np.zeros(0, dtype=np.int32)
            recv_dst = gathered[o].get(rank, np.zeros(0, dtype=np.int32))
            peers.append((o, send_idx, recv_dst))
    ctx.md_exchange_setup(gidx[mine], part["ghost_local"][mine], peers)
    part["exchange"] = {"own_ghosts": int(mine.sum()), "peers": [(o, int(len(a)), int(len(b))) for o, a, b in peers]}
    return ctx, cfg, part


def md_partition_check(torch, capi, synth, dist, device, rank, world, uid, cells=12):
    """Outside every timed region: a small box cut over the ranks must give, after the exchange, the forces and the
    energy of the same box on ONE rank (north_star: forces within 1e-10)."""
    ctx, cfg, part = md_partition(torch, capi, synth, dist, device, cells, rank, world, comm_uid=uid)
    x, nlocal = part["x"], part["nlocal"]
    f = np.zeros((nlocal, 3)); ea = np.zeros(max(nlocal, 1))
    xp = torch.from_numpy(x).pin_memory(); fp = torch.from_numpy(f).pin_memory(); ep = torch.from_numpy(ea).pin_memory()
    et = ctx.md_step_exchange_ptr(xp.data_ptr(), fp.data_ptr(), ep.data_ptr())
    ctx.close()
    res = [None] * world
    if world > 1:
        dist.all_gather_object(res, (part["owned_ids"], fp.numpy().copy(), et))
    else:
        res[0] = (part["owned_ids"], fp.numpy().copy(), et)
    out = None
    if rank == 0:
        one, cfg1, p1 = md_partition(torch, capi, synth, None, device, cells, 0, 1)
        f1 = torch.zeros(p1["nlocal"], 3, dtype=torch.float64).pin_memory()
        x1 = torch.from_numpy(p1["x"]).pin_memory()
        et1 = one.md_step_exchange_ptr(x1.data_ptr(), f1.data_ptr(), None)
        one.close()
        fall = np.zeros_like(f1.numpy())
        for ids, fr, _ in res:
            fall[ids] = fr
        fmax = float(np.max(np.abs(f1.numpy())))
        out = {"workload": "diamond Si %d^3 cells (%d atoms) cut into %d slabs vs one rank" % (cells, len(fall), world),
               "force_rel_err": float(np.max(np.abs(fall - f1.numpy())) / fmax), "energy_rel_err": abs(res[0][2] - et1) / abs(et1),
               "energies_equal_on_all_ranks": bool(all(abs(r[2] - res[0][2]) == 0.0 for r in res)), "tolerance": 1e-10}
        out["ok"] = bool(out["force_rel_err"] <= 1e-10 and out["energy_rel_err"] <= 1e-12 and out["energies_equal_on_all_ranks"])
    return out


def run_md_bench(torch, capi, synth, device, cells, steps, warmup, dist=None, rank=0, world=1):
    """pair_nn as BASELINE.json names it: ONE box (cells^3 diamond cells; C5 = 50 -> 1 M atoms) cut into `world` slabs
    along x, one per GPU (strong scaling).  pp_md_build_neighbors once (LAMMPS rebuilds every 10 steps, pair_nn.cpp:276),
    then `steps` calls of the PairNN::compute equivalent followed by the reverse communication of the ghost forces and the
    energy sum -- both owned by the library (pp_md_exchange_device: pack kernel, NCCL send/recv, ordered add) and both
    inside the timed regions: device-resident (value) and through host buffers (e2e)."""
    uid = None
    if world > 1:
        t = torch.tensor(list(capi.Context.comm_unique_id()) if rank == 0 else [0] * 128, dtype=torch.uint8, device="cuda")
        dist.broadcast(t, 0)
        uid = bytes(t.cpu().tolist())
    ctx, cfg, part = md_partition(torch, capi, synth, dist, device, cells, rank, world, comm_uid=uid)
    x, nlocal = part["x"], part["nlocal"]
    ntotal = len(x)
    stream = torch.cuda.ExternalStream(ctx.stream)
    dev = torch.device("cuda", device)
    xd = torch.from_numpy(x).to(dev)
    fd = torch.zeros_like(xd)
    ed = torch.zeros(1, dtype=torch.float64, device=dev)
    torch.cuda.synchronize()

    def one_step():
        with torch.cuda.stream(stream):
            fd.zero_()                                                    # LAMMPS clears forces before Pair::compute
        ctx.md_step_device(xd.data_ptr(), fd.data_ptr(), ed.data_ptr())
        ctx.md_exchange_device(fd.data_ptr(), ed.data_ptr())

    def barrier():
        torch.cuda.synchronize()
        if dist is not None:
            dist.barrier()
        torch.cuda.synchronize()

    for _ in range(warmup):
        one_step()
    barrier()
    ev0 = torch.cuda.Event(enable_timing=True); ev1 = torch.cuda.Event(enable_timing=True)
    if world > 1:
        one_step()                 # lines the ranks' timelines up after the host barrier (the exchange is a rendezvous)
    l0 = ctx.kernel_launches
    ev0.record(stream)
    for _ in range(steps):
        one_step()
    ev1.record(stream)
    barrier()
    dev_ms = ev0.elapsed_time(ev1) / steps
    launches = ctx.kernel_launches - l0
    exch_ms = ctx.phase_ms(7)
    # property check (outside the timed region): with every ghost force back on its owner the forces on the atoms of the
    # periodic box must add up to zero
    one_step()
    with torch.cuda.stream(stream):
        fsum_v = fd[:nlocal].sum(dim=0)
        fmax_v = fd[:nlocal].abs().max().reshape(1)
    torch.cuda.synchronize()
    if dist is not None:
        dist.all_reduce(fsum_v)
        dist.all_reduce(fmax_v, op=dist.ReduceOp.MAX)
    fsum = float(fsum_v.abs().max() / fmax_v[0])
    # e2e: pinned host positions in, forces of the local atoms / energies out, exchange included, every step
    x_pin = torch.from_numpy(x).pin_memory()
    f_pin = torch.zeros(max(nlocal, 1), 3, dtype=torch.float64).pin_memory()
    ea_pin = torch.zeros(max(nlocal, 1), dtype=torch.float64).pin_memory()
    for _ in range(max(1, warmup - 1)):
        ctx.md_step_exchange_ptr(x_pin.data_ptr(), f_pin.data_ptr(), ea_pin.data_ptr())
    barrier()
    if world > 1:
        ctx.md_step_exchange_ptr(x_pin.data_ptr(), f_pin.data_ptr(), ea_pin.data_ptr())      # lines the ranks up, as above
    phases = []
    t0 = time.perf_counter()
    for _ in range(steps):
        ctx.md_step_exchange_ptr(x_pin.data_ptr(), f_pin.data_ptr(), ea_pin.data_ptr())
        phases.append((ctx.phase_ms(4), ctx.phase_ms(5), ctx.phase_ms(6), ctx.phase_ms(7)))
    e2e_ms = (time.perf_counter() - t0) * 1e3 / steps
    ph = np.median(np.array(phases), axis=0)
    tt = torch.tensor([dev_ms, e2e_ms], dtype=torch.float64, device=dev)
    cnt = torch.tensor([float(nlocal), float(ntotal - nlocal)], dtype=torch.float64, device=dev)
    if dist is not None:
        dist.all_reduce(tt, op=dist.ReduceOp.MAX)
        dist.all_reduce(cnt)
    dev_ms, e2e_ms = [float(v) for v in tt.cpu()]
    atoms_total, ghosts_total = [int(v) for v in cnt.cpu()]
    out = {"metric": "pair_nn atom-steps/sec", "value": atoms_total / (dev_ms * 1e-3), "unit": "atom-steps/s", "ms_per_step": dev_ms,
           "n_gpus": world, "scaling": "strong", "atoms_total": atoms_total, "atoms_local": int(nlocal), "atoms_ghost": int(ntotal - nlocal),
           "ghosts_total": ghosts_total, "steps": steps, "warmup": warmup,
           "workload": "C5: ONE diamond-Si box of %dx%dx%d cells (%d atoms) cut into %d slab(s) along x, Rcut 6 + skin 1, 50 descriptors, "
                       "net 50-40-40-1" % (cells, cells, cells, atoms_total, world),
           "partition": "rank r owns x in [r L/N, (r+1) L/N); ghosts = periodic images within Rcut + skin; forces are computed on locals "
                        "and ghosts (pair_nn.cpp:182) and the ghost forces return to their owners through pp_md_exchange_device (pack "
                        "kernel, NCCL send/recv, ordered add; own periodic images folded locally), energy all-reduced -- inside both timings",
           "exchange": dict(part["exchange"], ms=exch_ms),
           "e2e": {"value": atoms_total / (e2e_ms * 1e-3), "unit": "atom-steps/s", "ms_per_step": e2e_ms,
                   "h2d_bytes_per_step": 24 * ntotal, "d2h_bytes_per_step": 24 * nlocal + 8 * nlocal + 8},
           "phases_ms": {"descriptors": float(ph[0]), "network": float(ph[1]), "forces": float(ph[2]), "exchange": float(ph[3])},
           "neighbor_build_s": part["build_s"], "gpu_launches": int(launches), "total_force_over_max_force": fsum}
    peaks = _fp64_peaks()
    # Analytic FP64 operation count of the factorised evaluation the kernels actually perform:
    # per neighbour and per unordered pair of neighbours, at C2's rows; SURVEY's row-by-row count is kept for reference.
    fl = md_flops_model(cfg, part.get("mean_neighbours", 45.2))
    out["roofline"] = {"bound": "fp64-fma", "kernel": "md_force_kernel", "achieved": fl["force"] * nlocal / (ph[2] * 1e-3) / 1e12,
                       "peak": peaks["fp64_fma_tflops"], "unit": "TFLOP/s",
                       "frac": fl["force"] * nlocal / (ph[2] * 1e-3) / 1e12 / peaks["fp64_fma_tflops"],
                       "algorithmic_flops_per_atom": fl["force"], "peak_source": peaks["source"], "flops_model": fl,
                       "descriptor_kernel": {"kernel": "md_desc_kernel", "algorithmic_flops_per_atom": fl["desc"],
                                             "achieved": fl["desc"] * nlocal / (ph[0] * 1e-3) / 1e12,
                                             "frac": fl["desc"] * nlocal / (ph[0] * 1e-3) / 1e12 / peaks["fp64_fma_tflops"]},
                       "naive_flops_per_atom": {"descriptors": 0.41e6, "forces": 1.21e6,
                                                "note": "SURVEY 8(d): every G4 row evaluated per ordered pair; not what runs"}}
    ctx.close()
    return out


def md_flops_model(cfg, n):
    """FP64 operations per central atom of the factorised power-basis evaluation (pp_desc.cuh), counting an FMA as two:
    n neighbours inside Rcut, n(n-1)/2 unordered pairs, NR radial keys (distinct eta), NB power levels x 2 lambdas."""
    NR = len({(r[0], r[1], r[2]) for r in cfg.G4})
    xis = sorted({int(r[3]) for r in cfg.G4})
    levels = len(xis)                                   # accumulated power levels (1, 2, 4, 16 -> 4)
    chain = int(np.log2(max(xis))) if xis else 0        # squarings from (1 + lam c) to the highest power
    nG2 = len(cfg.G2)
    pairs = n * (n - 1) / 2.0
    # per neighbour: distance, unit vector, cutoff spline + slope (~40), one exp + product per radial key and per G2 row
    per_nb_desc = 40 + 22 * NR + 24 * nG2
    # per unordered pair and lambda: cos (5), chain of squarings (chain), per level and radial key one FMA (2)
    per_pair_desc = 2 * (5 + 1 + chain + 2 * levels * NR) + 2 * NR
    # forces: the same products plus the derivative chain: per level and key two FMAs, three Cartesian components for each
    # of the two neighbours (prefix products of the powers give d/dc)
    per_nb_force = 60 + 30 * NR + 40 * nG2
    per_pair_force = 2 * (5 + 1 + 2 * chain + 4 * levels * NR) + 8 * NR + 36
    return {"neighbours": n, "radial_keys": NR, "power_levels": levels, "desc": n * per_nb_desc + pairs * per_pair_desc,
            "force": n * per_nb_force + pairs * per_pair_force,
            "note": "model of the factorised evaluation, FMA = 2; see DESIGN.md section 4.1"}


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
    key = {2: "mlp_fused2_kernel_%s_%d", 1: "mlp_kernel_fused_%s_%d", 3: "mlpw_kernels_%s_%d", 0: "two_phase_recompute_%s_%d",
           5: "two_phase_stored2_%s_%d", 4: "mlp32_fused_kernel_%s_%d"}.get(form, "form%d_%%s_%%d" % form) % (name, n_struct)
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
    ap.add_argument("--no-fp32", action="store_true", help="skip the fp32 object (FP32 mode on the headline workload, N = 1 only)")
    ap.add_argument("--md-cells", type=int, default=50, help="diamond-Si supercell edge in unit cells (C5 = 50 -> 1M atoms)")
    ap.add_argument("--md-steps", type=int, default=5)
    args = ap.parse_args()
    if args.impl == "reference":
        run_reference(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
