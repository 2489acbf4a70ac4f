#!/usr/bin/env python
"""bench.py — the headline benchmark of the suffix_sort path (BASELINE.json: k=256 SA build, Gsuffixes/s).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl reference] [--n BASES] [--k K]

One step = one pass of the whole path (pack -> get_lms -> k-ordered LMS sort -> put_lms -> induced_sort)
over one synthetic text.  At N=1 the workload is BASELINE.json configs[1]: uniform-random ACGT, 1 Gbp, k=256,
PARALLEL_SORTING.  `value` is device-resident throughput (text and SA in HBM, CUDA events on the library's
stream); `e2e` is the same call through the host C ABI with pinned host buffers (H2D of the text and D2H of
the suffix array inside the timed region).  At N>1 the LMS sort is sharded by key range with an NCCL
all-to-all; induction runs on rank 0.

`--impl reference` times the reference's own CPU implementation (oracle/_ref/libkiss_ref.so, the unmodified
headers compiled by oracle/build_ref.sh; else the C restatement) with all host threads on a bounded sample.
"""
import argparse
import ctypes as C
import json
import os
import statistics
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

N_DEFAULT = 1_000_000_000
SEED = 2


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--n", type=int, default=N_DEFAULT)
    ap.add_argument("--k", type=int, default=256)
    ap.add_argument("--cpu-sample", type=int, default=0, help="bases of the CPU baseline sample (0 = auto)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
    return ap.parse_args()


# ------------------------------------------------------------------------------------------ clocks
class ClockSampler:
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, index=0):
        self.rows = []
        self.proc = None
        self.index = index

    def start(self):
        try:
            self.proc = subprocess.Popen(
                ["nvidia-smi", f"--id={self.index}", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-lms", "100"],
                stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.th = threading.Thread(target=self._read, daemon=True)
            self.th.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append((time.time(), line.strip()))

    def window(self, t0, t1):
        self.t0, self.t1 = t0, t1

    def stop(self):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except Exception:
            self.proc.kill()
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        t0, t1 = getattr(self, "t0", 0.0), getattr(self, "t1", 1e18)
        inside = [r for (t, r) in self.rows if t0 <= t <= t1 + 0.15]
        for r in (inside or [r for (_, r) in self.rows]):
            f = [x.strip() for x in r.split(",")]
            if len(f) < 7:
                continue
            try:
                sm.append(float(f[0]))
                mx.append(float(f[1]))
            except ValueError:
                continue
            for nm, v in zip(names, f[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(nm)
        return {"sm_mhz": statistics.median(sm) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "samples": len(sm), "reasons": sorted(reasons)}


# ------------------------------------------------------------------------------------------ CPU reference
def cpu_reference_run(n, k, threads):
    """seconds of the reference's kiss1_suffix_array_dna on n bases of the bench text, with `threads` threads"""
    import numpy as np
    import oracle_lib as ol
    import textgen as tg
    S = tg.uniform(SEED, n)
    if ol.have_ref():
        L = ol.ref()
        SA = np.empty(n + 1, np.uint32)
        sec = L.ref_kiss1_suffix_array_dna(S.ctypes.data, n, k & 0xFFFFFFFF, threads, SA.ctypes.data)
        kind = "reference"
    else:
        t0 = time.perf_counter()
        SA = ol.oracle_kiss1(S, k)
        sec = time.perf_counter() - t0
        kind = "port"
    assert int(SA[0]) == n
    return sec, kind


def cpu_baseline(k, sample=0):
    threads = os.cpu_count() or 1
    if not sample:
        sec, kind = cpu_reference_run(20_000_000, k, threads)
        rate = 20_000_001 / sec
        sample = int(min(400_000_000, max(20_000_000, rate * 12)))
    sec, kind = cpu_reference_run(sample, k, threads)
    return {"value": (sample + 1) / sec / 1e9, "unit": "Gsuffixes/s", "cores": threads, "kind": kind,
            "sample": f"uniform ACGT seed {SEED}, first {sample} bases of the bench text, k={k}, "
                      f"kiss1_suffix_array_dna, {sec:.2f} s", "seconds": sec}


def run_reference_arm(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    threads = os.cpu_count() or 1
    sample = min(args.n, args.cpu_sample or 100_000_000)
    times = []
    kind = "reference"
    for i in range(args.warmup + args.steps):
        sec, kind = cpu_reference_run(sample, args.k, threads)
        if i >= args.warmup:
            times.append(sec)
    sec = sum(times) / len(times)
    val = (sample + 1) / sec / 1e9
    line = {
        "impl": "reference", "metric": "k=256 SA build throughput", "value": val, "unit": "Gsuffixes/s",
        "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup, "ms_per_step": sec * 1e3,
        "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "u32", "data": "synthetic",
        "config": {"workload": f"uniform-random ACGT {args.n} bp, k={args.k}, PARALLEL_SORTING "
                               f"(reference timed on a {sample}-base sample per step)"},
        "cpu_baseline": {"value": val, "unit": "Gsuffixes/s", "cores": threads, "kind": kind,
                         "sample": f"first {sample} bases of the bench text per step"},
        "e2e": {"value": val, "unit": "Gsuffixes/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)


# ------------------------------------------------------------------------------------------ B200 arm
# mean DRAM bytes (read + write) per launch at the 1 Gbp workload, ncu --set full (profiles/r1_final_launches_1gbp.txt)
NCU_DRAM_BYTES_PER_LAUNCH = {"k_wave_fused": 46.658e9 / 58, "k_radix_scatter": 5.4073e9, "k_groups_tiny": 6.7628e9}


def b200_alg_bytes(n, m):
    """B_alg(n, m) = 18 n + 28 m"""
    return 18 * n + 28 * m


def run_b200(args):
    import torch
    import kiss_b200
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dist = None
    if world > 1:
        import torch.distributed as dist
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    n, k = args.n, args.k
    ctx = kiss_b200.Context(local)
    lib_stream = torch.cuda.ExternalStream(kiss_b200.lib().kiss_b200_stream(ctx.handle), device=local)
    S = torch.empty(n + 64, dtype=torch.uint8, device="cuda")
    SA = torch.empty(n + 1, dtype=torch.int32, device="cuda")
    ctx.synth_uniform_dev(S.data_ptr(), n, SEED)
    torch.cuda.synchronize()

    if world > 1:
        from kiss_b200.multi_gpu import sharded_suffix_array_dev
        step = lambda: sharded_suffix_array_dev(ctx, S, n, SA, k, dist)
    else:
        step = lambda: ctx.suffix_array_dna_dev(S.data_ptr(), n, SA.data_ptr(), k)

    def barrier():
        if dist is not None:
            dist.barrier()
        torch.cuda.synchronize()

    clocks = ClockSampler(local)
    clocks.start()
    for _ in range(args.warmup):
        step()
    barrier()
    ctx.profile_enable(True)
    wall0 = time.time()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    stage = {}
    launches = 0
    e0.record(lib_stream)
    for _ in range(args.steps):
        t = step()
        launches += t["kernel_launches"]
        for key in ("pack", "get_lms", "lms_sort", "put_lms", "induce", "total_device", "exchange"):
            if key in t:
                stage[key] = stage.get(key, 0.0) + t[key]
        last = t
    e1.record(lib_stream)
    barrier()
    clocks.window(wall0, time.time())
    ms = e0.elapsed_time(e1)
    clk = clocks.stop()
    prof = ctx.profile()
    ctx.profile_enable(False)
    if dist is not None:
        tt = torch.tensor([ms], device="cuda")
        dist.all_reduce(tt, op=dist.ReduceOp.MAX)
        ms = float(tt.item())
    ms_step = ms / args.steps
    value = (n + 1) / (ms_step * 1e-3) / 1e9
    m = last["m"]

    line = {
        "metric": "k=256 SA build throughput", "value": value, "unit": "Gsuffixes/s", "n_gpus": world,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms_step, "higher_is_better": True,
        "scaling": "strong", "vs_baseline": None, "dtype": "u32", "data": "synthetic",
        "config": {"workload": f"uniform-random ACGT {n} bp (seed {SEED}), k={k}, PARALLEL_SORTING, "
                               f"{'LMS sort sharded over %d GPUs, induction on rank 0' % world if world > 1 else '1 B200'}",
                   "n": n, "m": m, "k": k, "l2": "inputs larger than L2 (text n bytes, SA 4(n+1) bytes)"},
        "clocks": clk,
        "gpu_launches": launches,
        "stages_ms": {kk: v / args.steps * 1e3 for kk, v in stage.items()},
        "waves": [last.get("waves_L"), last.get("waves_S")],
    }
    if "steps_ms" in last:
        line["sharded_sort_steps_ms"] = last["steps_ms"]
    if "shard_sizes" in last:
        line["shard_sizes"] = last["shard_sizes"]

    if rank == 0:
        # roofline of the dominant kernel class: algorithmic bytes per launch / its mean launch duration, measured
        # live with CUDA events around every launch (kiss_b200_profile_*), vs the measured copy bandwidth
        peaks = {}
        try:
            peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
        except Exception:
            pass
        peak = float(peaks.get("hbm_gbs", 6650.0))
        n_long = m - 1
        # one entry per kernel FUNCTION: the three profiler classes of k_wave_fused / k_wave_tail are one kernel (K5)
        merged = {}
        for name, (tot_ms, cnt) in prof.items():
            key = "k_wave_fused" if name.startswith("k_wave") else name
            a, b = merged.get(key, (0.0, 0))
            merged[key] = (a + tot_ms, b + cnt)
        steps = args.steps
        n_sort = n_long // world  # LMS suffixes one rank sorts (the sort shards, the rest runs on rank 0)
        alg = {  # algorithmic bytes per launch (SURVEY.md 8(d) per-unit figures x units of one launch)
            "k_radix_scatter": 2 * (4 + 4) * n_sort,            # read + write of (key32, pos32) per pass
            "k_radix_hist": 4 * n_sort,
            "k_groups_tiny": (4 + 4) * n_sort,
            "k_place/merge_lms": (4 + 4) * n_long,
            "k_pack_codes": n + n // 4,
        }
        if "k_wave_fused" in merged:
            # K5 induction (8(d)): 2 x 4(n+1) SA reads + 4n induced writes + 4m + 2 x n/4 packed text, per step,
            # spread over the launches of one step
            alg["k_wave_fused"] = (8 * (n + 1) + 4 * n + 4 * m + n // 2) / (merged["k_wave_fused"][1] / steps)
        kernels = {}
        for name, (tot_ms, cnt) in merged.items():
            kernels[name] = {"ms_per_step": tot_ms / steps, "launches_per_step": cnt / steps}
            if name in alg:
                gbs = alg[name] / (tot_ms / cnt * 1e-3) / 1e9
                kernels[name].update(alg_bytes_per_launch=alg[name], achieved_gbs=gbs, frac=gbs / peak)
        for name, (tot_ms, cnt) in prof.items():
            if name.startswith("k_wave"):
                kernels["k_wave_fused"].setdefault("classes", {})[name] = {"ms_per_step": tot_ms / steps,
                                                                          "launches_per_step": cnt / steps}
        line["kernels"] = kernels
        dom = max((k for k in kernels if "achieved_gbs" in kernels[k]), key=lambda k: kernels[k]["ms_per_step"])
        # dram__bytes_read.sum + dram__bytes_write.sum per launch (mean over the launches of one step) from the
        # ncu capture under profiles/ (r1_v6_dram_per_kernel.txt); None when the dominant kernel has no capture
        traffic = NCU_DRAM_BYTES_PER_LAUNCH.get(dom)
        line["roofline"] = {"bound": "hbm", "kernel": dom, "achieved": kernels[dom]["achieved_gbs"], "peak": peak,
                            "unit": "GB/s", "frac": kernels[dom]["frac"], "traffic": traffic,
                            "peak_source": "MEASURED_PEAKS.json hbm_gbs (measured copy)" if peaks else "fallback 6650 GB/s",
                            "share_of_step": kernels[dom]["ms_per_step"] / ms_step,
                            "launches_per_step": kernels[dom]["launches_per_step"]}
        if "lms_sort" in stage:  # the stage the north star asks to be reported at 2/4/8 GPUs as well
            sort_s = stage["lms_sort"] / args.steps
            line["lms_sort_stage"] = {"ms": sort_s * 1e3, "m": m, "value": m / sort_s / 1e9, "unit": "G LMS suffixes/s",
                                      "alg_bytes": 16 * m, "frac": 16 * m / sort_s / 1e9 / peak}
        balg = b200_alg_bytes(n, m)
        dev_s = stage.get("total_device", ms * 1e-3) / args.steps
        line["roofline_whole_path"] = {"alg_bytes": balg, "achieved": balg / dev_s / 1e9, "frac": balg / dev_s / 1e9 / peak,
                                       "note": "B_alg = 18 n + 28 m (SURVEY.md 8(d))"}

    # ---- end to end through the host C ABI (pinned host buffers, H2D + D2H inside the timed region)
    if world == 1 and not args.no_e2e:
        import numpy as np
        hS = torch.empty(n, dtype=torch.uint8).pin_memory()
        hSA = torch.empty(n + 1, dtype=torch.int32).pin_memory()
        hS.copy_(S[:n])
        torch.cuda.synchronize()
        del SA
        torch.cuda.empty_cache()
        hS_np = hS.numpy()
        hSA_np = hSA.numpy().view(np.uint32)
        ctx.suffix_array_dna(hS_np, k, out=hSA_np)  # warm-up
        t0 = time.perf_counter()
        reps = max(1, min(args.steps, 2))
        for _ in range(reps):
            ctx.suffix_array_dna(hS_np, k, out=hSA_np)
        torch.cuda.synchronize()
        sec = (time.perf_counter() - t0) / reps
        assert int(hSA_np[0]) == n
        line["e2e"] = {"value": (n + 1) / sec / 1e9, "unit": "Gsuffixes/s", "h2d_bytes_per_step": n,
                       "d2h_bytes_per_step": 4 * (n + 1), "ms_per_step": sec * 1e3,
                       "h2d_ms": ctx.last_timings["h2d"] * 1e3, "d2h_ms": ctx.last_timings["d2h"] * 1e3}
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        try:
            line["cpu_baseline"] = cpu_baseline(k, args.cpu_sample)
        except Exception as e:  # the checker is optional for the number, never for the product
            line["cpu_baseline"] = {"value": None, "unit": "Gsuffixes/s", "cores": os.cpu_count(), "kind": "unavailable",
                                    "sample": f"failed: {e}"}
    if rank == 0:
        print(json.dumps(line), flush=True)
    ctx.close()
    if dist is not None:
        dist.destroy_process_group()


def main():
    args = parse()
    if args.impl == "reference":
        run_reference_arm(args)
    else:
        run_b200(args)


if __name__ == "__main__":
    main()
