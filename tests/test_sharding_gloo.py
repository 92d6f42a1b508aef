"""The N > 1 path on CPU: two ranks (gloo) each take their contiguous shard of a batch, run the
oracle on it, and the gathered result equals the single-rank result; the max-over-ranks timing
reduction bench.py uses is exercised too.  The data path itself needs no collective."""
import os
import socket
import sys

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from conftest import ROOT


def shard_bounds(n, world, rank):
    """Same rule as b2i_predict_u8: GPU g takes [g*per, min(n, (g+1)*per)), per = ceil(n / G)."""
    per = (n + world - 1) // world
    return min(n, rank * per), min(n, (rank + 1) * per)


def _worker(rank, world, port, n, out_dir):
    sys.path.insert(0, ROOT)
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    from oracle import synth, tch_oracle
    torch.set_num_threads(2)
    imgs = synth.synthetic_images(n, seed=3)
    lo, hi = shard_bounds(n, world, rank)
    idx, prob, _ = tch_oracle.predict_u8("alexnet", synth.alexnet_weights(), imgs[lo:hi], k=5)
    gathered = [None] * world
    dist.all_gather_object(gathered, (lo, hi, idx, prob))
    t = torch.tensor([float(rank + 1)])
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    if rank == 0:
        full_idx = np.concatenate([g[2] for g in sorted(gathered, key=lambda g: g[0])])
        full_prob = np.concatenate([g[3] for g in sorted(gathered, key=lambda g: g[0])])
        np.savez(os.path.join(out_dir, "gathered.npz"), idx=full_idx, prob=full_prob, tmax=t.numpy(),
                 bounds=np.array([(g[0], g[1]) for g in gathered]))
    dist.barrier()
    dist.destroy_process_group()


def test_shard_bounds_cover_every_image_once():
    for n in (0, 1, 5, 8, 255, 256, 257, 1000):
        for world in (1, 2, 4, 8):
            spans = [shard_bounds(n, world, r) for r in range(world)]
            assert spans[0][0] == 0 and spans[-1][1] == n
            assert all(a[1] == b[0] for a, b in zip(spans, spans[1:]))


def test_two_rank_sharded_run_equals_single_rank(tmp_path):
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        port = s.getsockname()[1]
    n = 5  # odd: ragged shards (3 + 2)
    mp.spawn(_worker, args=(2, port, n, str(tmp_path)), nprocs=2, join=True)
    got = np.load(os.path.join(tmp_path, "gathered.npz"))
    sys.path.insert(0, ROOT)
    from oracle import synth, tch_oracle
    idx, prob, _ = tch_oracle.predict_u8("alexnet", synth.alexnet_weights(), synth.synthetic_images(n, seed=3), k=5)
    assert np.array_equal(got["idx"], idx)
    assert np.allclose(got["prob"], prob, rtol=0, atol=1e-6)
    assert got["tmax"][0] == 2.0
    assert sorted(map(tuple, got["bounds"])) == [(0, 3), (3, 5)]
