"""world_size-2 tests of the N > 1 host-side logic on CPU (gloo): structures are dealt round robin over ranks
(Setup.cpp:582), each rank forms [gradient, SSE, vSSE, nval] on its shard, and ONE sum over ranks -- the
exchange pp_epoch performs with NCCL on the GPUs -- gives what a single rank computes on all structures.
The per-rank arithmetic here is the oracle's (no GPU in this container); the GPU path's N = 2 run is bench.py."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

import oracle
from golden_util import TrainCase


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, name, out_dir):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    c = TrainCase(name)
    mine = list(range(rank, c.n, world))                       # Setup::split_systems
    off = np.zeros(len(mine) + 1, dtype=np.int32)
    off[1:] = np.cumsum([c.natoms[s] for s in mine])
    G = np.concatenate([c.G[s].T for s in mine], axis=0)
    elem = np.concatenate([c.elem[s] for s in mine])
    params = c.params_at(0)
    nets = [oracle.Net(c.sizes, params[e], c.mean(z), c.scale(z), node_transfer=c.node_transfer(z)) for e, z in enumerate(c.elements)]
    r = oracle.epoch(off, elem, G, c.targets[mine], c.tags[mine], nets, c.out_mean, c.out_scale)
    buf = torch.from_numpy(np.concatenate([r["gradient"], [r["SSE"], r["vSSE"], float(r["nval"])]]))
    dist.all_reduce(buf)                                        # the one exchange per epoch
    if rank == 0:
        np.save(os.path.join(out_dir, "reduced.npy"), buf.numpy())
    # the host optimizer runs redundantly on every rank from the identical reduced buffer: identical next parameters
    from prophet_b200 import hostapi
    opt = hostapi.Optimizer(c.d["x_before.0"], "rprop", int(c.d["Ntrain"][0]))
    x1 = opt.step(float(buf[-3]), buf[:-3].numpy())
    xs = [torch.zeros(len(x1), dtype=torch.float64) for _ in range(world)]
    dist.all_gather(xs, torch.from_numpy(x1))
    assert all(torch.equal(xs[0], t) for t in xs)
    if rank == 0:
        np.save(os.path.join(out_dir, "x1.npy"), x1)
    dist.destroy_process_group()


@pytest.mark.parametrize("name", ["tio2_train", "si_train"])
def test_two_ranks_reduce_to_single_rank_result(name, tmp_path):
    port = _free_port()
    mp.spawn(_worker, args=(2, port, name, str(tmp_path)), nprocs=2, join=True)
    c = TrainCase(name)
    red = np.load(str(tmp_path / "reduced.npy"))
    g_ref = c.d["gradient.0"]
    assert np.max(np.abs(red[:-3] - g_ref)) <= 1e-10 * np.max(np.abs(g_ref))
    assert abs(red[-3] - c.d["SSE.0"][0]) <= 1e-10 * abs(c.d["SSE.0"][0])
    assert abs(red[-2] - c.d["vSSE.0"][0]) <= 1e-10 * max(abs(c.d["vSSE.0"][0]), 1e-300)
    assert int(round(red[-1])) == int(c.d["nval.0"][0])
    # first Rprop step moves every parameter by -1e-4 regardless of the gradient (Optimizer.cpp:549-554)
    assert np.array_equal(np.load(str(tmp_path / "x1.npy")), c.d["x_after.0"])


# ---------------------------------------------------------------------------------------------------------------
# pair_nn on a box cut into slabs (the C5 benchmark at N > 1): host-side logic of the partition and of the ghost-force
# exchange lists, with the oracle's force routine standing in for the GPU kernels and gloo for NCCL.
def _md_worker(rank, world, port, out_dir):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    from prophet_b200 import synth
    cfg = synth.config_c5(3)                       # 216 atoms, box 16.3 A: slabs of 8.1 A, halo 7 A
    cfg.G4 = cfg.G4[:4]
    cell, cart, Z, _, _ = synth.make_structure(cfg, 5)
    halo = cfg.rcut + 1.0
    part = synth.slab_partition(cell, cart, rank, world, halo)
    x, nlocal = part["x"], part["nlocal"]
    numneigh, offsets, flat = synth.brute_full_neighbor_list(x, nlocal, halo)
    rng = np.random.RandomState(7)
    dEdG_all = rng.standard_normal((len(cart), cfg.NG))          # the same "network" derivative for every global atom
    dEdG = dEdG_all[part["owned_ids"]]
    f = oracle.forces_md(x, np.zeros(len(x), dtype=np.int32), nlocal, offsets, flat, 1, cfg.G1, cfg.G2, cfg.G3, cfg.G4, cfg.rcut, dEdG)
    # reverse communication exactly as bench.py sets it up for pp_md_exchange_setup
    gidx = np.arange(nlocal, len(x))
    mine = part["ghost_rank"] == rank
    np.add.at(f, part["ghost_local"][mine], f[gidx[mine]])
    wanted = {int(o): (gidx[part["ghost_rank"] == o], part["ghost_local"][part["ghost_rank"] == o])
              for o in np.unique(part["ghost_rank"]) if int(o) != rank}
    gathered = [None] * world
    dist.all_gather_object(gathered, {o: (v[1], f[v[0]]) for o, v in wanted.items()})
    for r in range(world):
        if r != rank and rank in gathered[r]:
            dst, vals = gathered[r][rank]
            np.add.at(f, dst, vals)
    res = [None] * world
    dist.all_gather_object(res, (part["owned_ids"], f[:nlocal]))
    if rank == 0:
        fall = np.zeros((len(cart), 3))
        for ids, fr in res:
            fall[ids] = fr
        np.save(os.path.join(out_dir, "f_partitioned.npy"), fall)
    dist.destroy_process_group()


def test_two_slabs_give_the_single_box_forces(tmp_path):
    from prophet_b200 import synth
    port = _free_port()
    mp.spawn(_md_worker, args=(2, port, str(tmp_path)), nprocs=2, join=True)
    cfg = synth.config_c5(3)
    cfg.G4 = cfg.G4[:4]
    cell, cart, Z, _, _ = synth.make_structure(cfg, 5)
    halo = cfg.rcut + 1.0
    one = synth.slab_partition(cell, cart, 0, 1, halo)
    assert one["nlocal"] == len(cart) and np.all(one["ghost_rank"] == 0)
    numneigh, offsets, flat = synth.brute_full_neighbor_list(one["x"], one["nlocal"], halo)
    rng = np.random.RandomState(7)
    dEdG_all = rng.standard_normal((len(cart), cfg.NG))
    f = oracle.forces_md(one["x"], np.zeros(len(one["x"]), dtype=np.int32), one["nlocal"], offsets, flat, 1, cfg.G1, cfg.G2, cfg.G3,
                         cfg.G4, cfg.rcut, dEdG_all)
    np.add.at(f, one["ghost_local"], f[one["nlocal"]:])
    f_one = f[:one["nlocal"]]
    f_two = np.load(str(tmp_path / "f_partitioned.npy"))
    assert np.max(np.abs(f_two - f_one)) <= 1e-10 * np.max(np.abs(f_one))
    assert np.max(np.abs(f_one.sum(axis=0))) <= 1e-9 * np.max(np.abs(f_one))       # a periodic box: no net force
