"""Sharded LMS sort: one process per GPU, torch.distributed (NCCL over NVLink) for the plumbing.

Only the k-ordered LMS sort shards (BASELINE.json north_star): the text is replicated, every
rank classifies it, rank r owns the r-th slice of the text-ordered LMS list, the slices are bucketed by
sampled key splitters, ONE personalised all-to-all moves LMS positions (4 bytes each; keys are re-gathered
from the replicated text), every rank sorts its key range with the single-GPU kernels, and the sorted ranges
are concatenated on rank 0, where placement and induction (a single dependency chain) run.

Equal first-16-base keys always share a destination rank, so the tie refinement and the position tie-break
stay local and the result is bit-identical to the single-GPU path.

The orchestration is written against a small backend interface so that the same code runs under gloo on the
CPU in tests/test_multi_gpu_host.py (backend = numpy + the oracle) and under NCCL on GPUs (backend = the C ABI).
"""
import numpy as np

SAMPLES_PER_RANK = 4096


def choose_splitters(all_samples, world):
    """world-1 ascending splitters from the concatenated key samples of all ranks (uint32).
    destination of a key = number of splitters <= key."""
    s = np.sort(np.asarray(all_samples, dtype=np.uint32))
    if s.size == 0 or world <= 1:
        return np.zeros(0, np.uint32)
    idx = (np.arange(1, world, dtype=np.int64) * s.size) // world
    return s[idx].astype(np.uint32)


def slice_bounds(n_long, rank, world):
    return (n_long * rank) // world, (n_long * (rank + 1)) // world


class GpuBackend:
    """device arrays are torch int32 tensors on the rank's GPU; kernels via the C ABI"""

    def __init__(self, ctx, torch):
        self.ctx, self.torch = ctx, torch

    def prepare(self, S, n, k):
        self.m = self.ctx.prepare_dev(S.data_ptr(), n, k)
        self.kp, self.pp, self.n_long = self.ctx.lms_keys_dev()
        return self.n_long

    def sample_keys(self, lo, hi, count):
        torch = self.torch
        if hi <= lo:
            return np.zeros(0, np.uint32)
        # strided sample of the slice's keys, read straight from the context's device array
        cnt = min(count, hi - lo)
        idx = (np.arange(cnt, dtype=np.int64) * (hi - lo)) // cnt + lo
        keys = _as_tensor(torch, self.kp, self.n_long)
        return keys[torch.from_numpy(idx).cuda()].cpu().numpy().view(np.uint32)

    def partition(self, splitters, lo, hi):
        torch = self.torch
        out = torch.empty(max(hi - lo, 1), dtype=torch.int32, device="cuda")
        counts = self.ctx.partition_lms_dev(splitters, lo, hi, out.data_ptr())
        return out[:hi - lo], counts.astype(np.int64)

    def empty(self, count):
        return self.torch.empty(max(count, 1), dtype=self.torch.int32, device="cuda")[:count]

    def sort_subset(self, buf, count):
        self.torch.cuda.synchronize()
        return self.ctx.sort_lms_subset_dev(buf.data_ptr(), count)

    def contexts(self, buf, count):
        """induction contexts of the sorted suffixes buf[0..count) (one random text fetch each: this is the part of
        the induction's first touch of an LMS suffix that shards with the sort)"""
        out = self.empty(count)
        self.torch.cuda.synchronize()
        self.ctx.lms_contexts_dev(buf.data_ptr(), count, out.data_ptr())  # enqueued on the library's stream
        return out

    def contexts_wait(self):
        import kiss_b200
        self.torch.cuda.ExternalStream(kiss_b200.lib().kiss_b200_stream(self.ctx.handle)).synchronize()

    def finish(self, sorted_long, n_long, SA, ctx_long=None):
        self.torch.cuda.synchronize()
        return self.ctx.finish_from_sorted_dev(sorted_long.data_ptr(), n_long, SA.data_ptr(),
                                               ctx_long.data_ptr() if ctx_long is not None and n_long else None)

    def sync(self):
        self.torch.cuda.synchronize()


class PeerBuffers:
    """(position, context) buffers of n_long words in symmetric memory: every rank's copy is mapped into every process
    (CUDA peer memory over NVLink / NVSwitch), so a kernel or a copy on rank r stores straight into rank 0's buffer.
    Creation is collective; `ok` is False on every rank when any rank could not map the buffers."""

    def __init__(self, torch, dist, n_long):
        self.cap, self.ok = n_long, False

        def agree(flag):  # every rank learns whether ALL ranks got this far (a collective every rank always enters)
            t = torch.tensor([1 if flag else 0], dtype=torch.int32, device="cuda")
            dist.all_reduce(t, op=dist.ReduceOp.MIN)
            return bool(int(t.item()))

        # 1. the module and the local allocations: nothing collective yet, so a rank that fails here cannot leave the
        # others waiting inside a rendezvous
        symm = None
        try:
            import torch.distributed._symmetric_memory as symm
            dev = torch.device("cuda", torch.cuda.current_device())
            self.pos = symm.empty(max(n_long, 4), dtype=torch.int32, device=dev)
            self.ctx = symm.empty(max(n_long, 4), dtype=torch.int32, device=dev)
        except Exception as e:  # no symmetric memory on this system: the NCCL concatenation is used
            self.err = repr(e)
            symm = None
        if not agree(symm is not None):
            return
        # 2. the rendezvous is collective and every rank enters it; its outcome is agreed on again
        good = True
        try:
            self.hpos = symm.rendezvous(self.pos, dist.group.WORLD)
            self.hctx = symm.rendezvous(self.ctx, dist.group.WORLD)
            self.pos0 = self.hpos.get_buffer(0, (max(n_long, 4),), torch.int32)
            self.ctx0 = self.hctx.get_buffer(0, (max(n_long, 4),), torch.int32)
        except Exception as e:
            self.err = repr(e)
            good = False
        self.ok = agree(good)


def peer_buffers(backend, dist, n_long):
    """cached on the library context (bench steps repeat the same text length)"""
    import os
    if os.environ.get("KISS_B200_NO_PEER"):
        return None
    pb = getattr(backend.ctx, "_peer_buffers", None)
    if pb is None or pb.cap != n_long:
        pb = PeerBuffers(backend.torch, dist, n_long)
        backend.ctx._peer_buffers = pb
    return pb if pb.ok else None


def _as_tensor(torch, ptr, count):
    """int32 view of `count` device words at raw address `ptr` (no copy)"""
    class _Arr:
        __cuda_array_interface__ = {"shape": (count,), "typestr": "<i4", "data": (ptr, False), "version": 2}
    return torch.as_tensor(_Arr(), device="cuda")


def sharded_lms_sort(backend, dist, n_long, rank, world):
    """returns (sorted_long on rank 0 else None, stats).  Collectives: one all_gather of key samples, one
    all_to_all of counts, one all_to_all of positions, one gather-shaped all_to_all of the sorted ranges."""
    torch = backend.torch
    import os
    import time
    marks = []

    def mark(name):  # KISS_B200_MGPU_TIMING=1: synchronised wall-clock marks between the steps (debug aid)
        if os.environ.get("KISS_B200_MGPU_TIMING"):
            backend.sync()
            marks.append((name, time.perf_counter()))

    mark("start")
    lo, hi = slice_bounds(n_long, rank, world)
    # 1. splitters
    mine = backend.sample_keys(lo, hi, SAMPLES_PER_RANK)
    pad = np.full(SAMPLES_PER_RANK + 1, 0, np.int64)
    pad[0] = mine.size
    pad[1:1 + mine.size] = mine
    t = torch.from_numpy(pad).to(backend_device(backend))
    gathered = [torch.empty_like(t) for _ in range(world)]
    dist.all_gather(gathered, t)
    samples = np.concatenate([g.cpu().numpy()[1:1 + int(g[0].item())] for g in gathered]).astype(np.uint32)
    splitters = choose_splitters(samples, world)
    mark("splitters")
    # 2. bucket this rank's slice by destination (stable) and exchange the counts
    send, send_counts = backend.partition(splitters, lo, hi)
    sc = torch.from_numpy(send_counts.astype(np.int64)).to(backend_device(backend))
    rc = torch.empty_like(sc)
    dist.all_to_all_single(rc, sc)
    recv_counts = rc.cpu().numpy()
    total = int(recv_counts.sum())
    mark("partition")
    # 3. the exchange: LMS positions only
    recv = backend.empty(total)
    backend.sync()
    dist.all_to_all_single(recv, send, output_split_sizes=recv_counts.tolist(), input_split_sizes=send_counts.tolist())
    mark("exchange")
    # chunks arrive in source-rank order = ascending text order: the stable sort's precondition
    # 4. local sort of this rank's key range
    st = backend.sort_subset(recv, total)
    mark("local_sort")
    # 5. concatenate the ranges on rank 0 (key ranges are ordered by rank)
    sizes = torch.tensor([total], dtype=torch.int64, device=backend_device(backend))
    all_sizes = [torch.empty_like(sizes) for _ in range(world)]
    dist.all_gather(all_sizes, sizes)
    all_sizes = [int(x.item()) for x in all_sizes]
    pb = peer_buffers(backend, dist, n_long) if hasattr(backend, "contexts") else None
    if pb is not None:
        # 5a. peer memory: the context kernel of rank r — one random text fetch per sorted suffix, the part of the
        # induction's first touch of an LMS suffix that shards with the sort — stores its results straight into rank
        # 0's context buffer over NVLink, and the sorted positions follow with one device-to-device copy into rank
        # 0's position buffer, both at this rank's offset in the global order.  No staging, no NCCL on this step.
        off = sum(all_sizes[:rank])
        backend.sync()
        if total:
            backend.ctx.lms_contexts_dev(recv.data_ptr(), total, pb.ctx0.data_ptr() + 4 * off)  # library stream
            pb.pos0[off:off + total].copy_(recv)                                                # torch stream
        backend.sync()   # both streams: this rank's stores have left
        dist.barrier()   # ... and everybody else's have landed
        mark("concat_peer")
        stats = {"shard_sizes": all_sizes, "imbalance": (max(all_sizes) * world / max(1, sum(all_sizes))), "sort": st,
                 "ctx": pb.ctx[:n_long] if rank == 0 else None, "concat": "peer memory"}
        if marks:
            stats["steps_ms"] = {marks[i][0]: round((marks[i][1] - marks[i - 1][1]) * 1e3, 3) for i in range(1, len(marks))}
        return (pb.pos[:n_long] if rank == 0 else None), stats
    # 5b. NCCL: contexts fetched next to the exchange of the positions, then a second gather-shaped all_to_all
    ctx_local = backend.contexts(recv, total) if hasattr(backend, "contexts") else None
    mark("contexts_enqueued")
    out_split = all_sizes if rank == 0 else [0] * world
    in_split = [total] + [0] * (world - 1)
    result = backend.empty(sum(out_split))
    backend.sync()
    dist.all_to_all_single(result, recv, output_split_sizes=out_split, input_split_sizes=in_split)
    ctx_result = None
    if ctx_local is not None:  # (the context kernel ran next to the exchange of the positions)
        backend.contexts_wait()
        ctx_result = backend.empty(sum(out_split))
        dist.all_to_all_single(ctx_result, ctx_local, output_split_sizes=out_split, input_split_sizes=in_split)
    mark("concat")
    stats = {"shard_sizes": all_sizes, "imbalance": (max(all_sizes) * world / max(1, sum(all_sizes))),
             "sort": st, "ctx": ctx_result if rank == 0 else None, "concat": "nccl"}
    if marks:
        stats["steps_ms"] = {marks[i][0]: round((marks[i][1] - marks[i - 1][1]) * 1e3, 3) for i in range(1, len(marks))}
    return (result if rank == 0 else None), stats


def backend_device(backend):
    return getattr(backend, "device", "cuda")


def sharded_suffix_array_dev(ctx, S, n, SA, k, dist):
    """bench / product entry at N > 1: S, SA are torch tensors on this rank's GPU (S replicated on every rank);
    the suffix array lands in SA on rank 0.  Returns a timing dict shaped like Context.last_timings."""
    import time
    import torch
    rank, world = dist.get_rank(), dist.get_world_size()
    import kiss_b200
    be = GpuBackend(ctx, torch)
    launches0 = kiss_b200.lib().kiss_b200_launch_count()
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    n_long = be.prepare(S, n, k)
    t1 = time.perf_counter()
    sorted_long, stats = sharded_lms_sort(be, dist, n_long, rank, world)
    torch.cuda.synchronize()
    t2 = time.perf_counter()
    out = {"kernel_launches": 0, "m": be.m, "get_lms": t1 - t0, "lms_sort": t2 - t1, "put_lms": 0.0, "induce": 0.0,
           "pack": 0.0, "shard_sizes": stats["shard_sizes"], "imbalance": stats["imbalance"]}
    if "steps_ms" in stats:
        out["steps_ms"] = stats["steps_ms"]
    out["concat"] = stats.get("concat")
    if rank == 0:
        ft = be.finish(sorted_long, n_long, SA, stats.get("ctx"))
        out.update(put_lms=ft["put_lms"], induce=ft["induce"], waves_L=ft["waves_L"], waves_S=ft["waves_S"])
    # this rank's own kernels, counted by the library (prepare, partition, subset sort, contexts, finish)
    out["kernel_launches"] = int(kiss_b200.lib().kiss_b200_launch_count() - launches0)
    out["total_device"] = out["get_lms"] + out["lms_sort"] + out["put_lms"] + out["induce"]
    dist.barrier()
    return out
