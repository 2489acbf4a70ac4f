"""Host-side logic of the sharded LMS sort (kiss_b200/multi_gpu.py) under gloo, world_size 2, on the CPU:
splitter choice, stable bucketing, the all-to-all plan and the rank-ordered concatenation.  The device kernels
are replaced by a numpy + oracle backend; the orchestration code is the one the GPUs run."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

import oracle_lib as ol
import textgen as tg
from kiss_b200 import multi_gpu as mg

K = 32
D = 125 * (K // 125 + 1)


class NumpyBackend:
    device = "cpu"

    def __init__(self, S):
        self.torch = torch
        self.S = S
        n = S.size
        o = ol.oracle_kiss1(S, k=K, stages=True)
        lms = o["lms"][:-1].astype(np.int64)
        self.longs = lms[lms + D <= n]
        self.n_long = self.longs.size
        pad = np.concatenate([S, np.zeros(16, np.uint8)]).astype(np.uint64)
        key = np.zeros(self.n_long, np.uint64)
        for t in range(16):
            key = (key << np.uint64(2)) | pad[self.longs + t]
        self.keys = key.astype(np.uint32)
        rank = np.full(n + 1, -1, np.int64)
        rank[o["sorted"].astype(np.int64)] = np.arange(o["sorted"].size)
        self.rank_of = rank
        self.sorted_longs = np.array([p for p in o["sorted"][1:].astype(np.int64) if p + D <= n], np.int64)

    def sample_keys(self, lo, hi, count):
        cnt = min(count, hi - lo)
        if cnt <= 0:
            return np.zeros(0, np.uint32)
        idx = (np.arange(cnt, dtype=np.int64) * (hi - lo)) // cnt + lo
        return self.keys[idx]

    def partition(self, splitters, lo, hi):
        dest = np.searchsorted(splitters, self.keys[lo:hi], side="right")
        order = np.argsort(dest, kind="stable")
        counts = np.bincount(dest, minlength=splitters.size + 1).astype(np.int64)
        return torch.from_numpy(self.longs[lo:hi][order].astype(np.int32)), counts

    def empty(self, count):
        return torch.empty(count, dtype=torch.int32)

    def sort_subset(self, buf, count):
        a = buf.numpy().astype(np.int64)
        assert np.all(np.diff(a) > 0), "received positions must be in ascending text order (stability)"
        buf.copy_(torch.from_numpy(a[np.argsort(self.rank_of[a], kind="stable")].astype(np.int32)))
        return {"kernel_launches": 0}

    def sync(self):
        pass


def _worker(rank, world, port, seed, q):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        S = tg.planted_repeats(seed, 60_000, unit_len=1500, copies=20)
        be = NumpyBackend(S)
        res, stats = mg.sharded_lms_sort(be, dist, be.n_long, rank, world)
        if rank == 0:
            ok = np.array_equal(res.numpy().astype(np.int64), be.sorted_longs)
            q.put((ok, stats["shard_sizes"]))
    finally:
        dist.destroy_process_group()


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


@pytest.mark.parametrize("world", [2, 3])
def test_sharded_lms_sort_host_logic(world):
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, world, port, 77, q)) for r in range(world)]
    for p in procs:
        p.start()
    for p in procs:
        p.join(120)
        assert p.exitcode == 0
    ok, sizes = q.get(timeout=10)
    assert ok
    assert sum(sizes) > 0 and len(sizes) == world


def test_choose_splitters_keeps_equal_keys_together():
    keys = np.array([5] * 100 + [9] * 100 + [1], np.uint32)
    spl = mg.choose_splitters(keys, 4)
    dest = np.searchsorted(spl, keys, side="right")
    for v in (1, 5, 9):
        assert np.unique(dest[keys == v]).size == 1
    assert np.all(np.diff(spl.astype(np.int64)) >= 0)


def test_slice_bounds_cover_everything():
    for n in (0, 1, 7, 1000):
        for w in (1, 2, 8):
            b = [mg.slice_bounds(n, r, w) for r in range(w)]
            assert b[0][0] == 0 and b[-1][1] == n
            assert all(b[i][1] == b[i + 1][0] for i in range(w - 1))
