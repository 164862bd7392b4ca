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
