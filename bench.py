#!/usr/bin/env python
"""bench.py — the headline benchmark of the suffix_sort path (BASELINE.json: k=256 SA build, Gsuffixes/s).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl reference] [--n BASES] [--k K]

One step = one pass of the whole path (pack -> get_lms -> k-ordered LMS sort -> put_lms -> induced_sort)
over one synthetic text.  The workload is the one BASELINE.json's metric is quoted on ("k=256 SA build ... on
3.1 Gbp"): configs[2] = C3 of SURVEY.md 8(d), chm13-sized 3 117 292 070 bp with injected interspersed / tandem /
segmental repeats (tests/textgen.py chm13_like, seed 3), k=256, PARALLEL_SORTING; it fits one B200 (45 of
180 GB).  `--workload c2` selects configs[1] (uniform-random 1 Gbp, seed 2).  `value` is device-resident throughput (text and SA in HBM, CUDA events on the library's
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

# C2 = uniform 1 Gbp seed 2; C3 = chm13-sized with repeats, seed 3 (README.md:101 gives the length)
WORKLOADS = {
    "c2": dict(n=1_000_000_000, seed=2, kind="uniform", name="uniform-random ACGT"),
    "c3": dict(n=3_117_292_070, seed=3, kind="chm13", name="chm13-sized ACGT with injected tandem/interspersed/segmental repeats"),
}


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--workload", default="c3", choices=sorted(WORKLOADS))
    ap.add_argument("--n", type=int, default=0, help="bases (0 = the workload's own size)")
    ap.add_argument("--k", type=int, default=256)
    ap.add_argument("--cpu-sample", type=int, default=0, help="bases of the CPU baseline sample (0 = auto)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-full-config", action="store_true", help="reference arm: skip the one full-size run")
    ap.add_argument("--no-parity", action="store_true")
    args = ap.parse_args()
    args.wl = WORKLOADS[args.workload]
    args.n = args.n or args.wl["n"]
    return args


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
CPU_SAMPLE = 100_000_000  # bases of the bounded CPU sample: the SAME text for the cpu_baseline leg and the reference arm


def host_text(wl, n):
    """the workload's text at n bases, generated on the host with numpy only (the reference arm never touches the GPU);
    n = the workload's own size gives the bench text itself"""
    import numpy as np
    import textgen as tg
    S = tg.uniform_big(wl["seed"], n)
    if wl["kind"] != "uniform":
        S = tg.chm13_like(wl["seed"], n, base=S)
    return S


def host_sample(wl, sample):
    what = (f"first {sample} bases of the bench text" if wl["kind"] == "uniform" else
            f"the bench text's generator (chm13_like, seed {wl['seed']}) at {sample} bases")
    return host_text(wl, sample), what


def cpu_reference_run(S, k, threads):
    """seconds of the reference's kiss1_suffix_array_dna on the text S with `threads` threads"""
    import numpy as np
    import oracle_lib as ol
    n = S.size
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


def cpu_baseline(wl, k, sample=0):
    """the reference on the bounded sample (the same one the reference arm times), all host threads, best of 2"""
    threads = os.cpu_count() or 1
    sample = min(sample or CPU_SAMPLE, wl["n"])
    S, what = host_sample(wl, sample)
    runs = [cpu_reference_run(S, k, threads) for _ in range(2)]
    sec, kind = min(r[0] for r in runs), runs[0][1]
    return {"value": (sample + 1) / sec / 1e9, "unit": "Gsuffixes/s", "cores": threads, "kind": kind,
            "sample": f"{what} ({wl['name']}, seed {wl['seed']}), k={k}, kiss1_suffix_array_dna, best of 2: {sec:.2f} s",
            "seconds": sec}


def run_reference_arm(args):
    """K bounded-sample steps (the line's value), then — unless --no-full-config — ONE run of the full configured text,
    whose seconds go into `full_config` (the same-config number; 3.1 Gbp takes the reference about half a minute on 16
    threads), and one run at 32 threads where the host has them."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    threads = os.cpu_count() or 1
    wl = args.wl
    sample = min(args.n, args.cpu_sample or CPU_SAMPLE)
    S, what = host_sample(wl, sample)
    times = []
    kind = "reference"
    for i in range(args.warmup + args.steps):
        sec, kind = cpu_reference_run(S, args.k, threads)
        if i >= args.warmup:
            times.append(sec)
    sec = sum(times) / len(times)
    val = (sample + 1) / sec / 1e9
    line = {
        "impl": "reference", "metric": "k=256 SA build throughput", "value": val, "unit": "Gsuffixes/s",
        "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup, "ms_per_step": sec * 1e3,
        "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "u32", "data": "synthetic",
        "config": {"workload": f"{wl['name']} {args.n} bp (seed {wl['seed']}), k={args.k}, PARALLEL_SORTING "
                               f"(reference timed on a {sample}-base sample per step: {what}; one full-size run in full_config)"},
        "cpu_baseline": {"value": val, "unit": "Gsuffixes/s", "cores": threads, "kind": kind,
                         "sample": f"{what}, every step"},
        "e2e": {"value": val, "unit": "Gsuffixes/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    if threads >= 32:
        s32, _ = cpu_reference_run(S, args.k, 32)
        line["cpu_baseline"]["threads32"] = {"value": (sample + 1) / s32 / 1e9, "seconds": s32, "sample": "the same sample"}
    del S
    if not args.no_full_config and args.n > sample:
        t0 = time.perf_counter()
        F = host_text(wl, args.n)
        gen_s = time.perf_counter() - t0
        fsec, _ = cpu_reference_run(F, args.k, threads)
        full = {"n": args.n, "seconds": fsec, "value": (args.n + 1) / fsec / 1e9, "unit": "Gsuffixes/s", "cores": threads,
                "runs": 1, "text_generation_seconds": round(gen_s, 1),
                "note": "the whole configured text, kiss1_suffix_array_dna, same call the steps time"}
        line["full_config"] = full
        line["cpu_baseline"]["full_config_seconds"] = fsec
        line["cpu_baseline"]["full_config_value"] = full["value"]
    print(json.dumps(line), flush=True)


# ------------------------------------------------------------------------------------------ B200 arm
def ncu_traffic(workload, kernel):
    """mean dram__bytes_read.sum + dram__bytes_write.sum per launch of `kernel` from the newest ncu launch list of this
    workload: profiles/ncu_dram_per_launch.json, regenerated by tools/profile_c3.sh (tools/ncu_dram_per_kernel.py);
    None when there is no capture"""
    try:
        d = json.load(open(os.path.join(ROOT, "profiles", "ncu_dram_per_launch.json")))[workload]
    except Exception:
        return None, None
    per = d["dram_bytes_per_launch"]
    names = {"k_wave_fused": ["k_wave_fused", "k_wave_tail"], "k_radix_scatter": ["k_radix_scatter_p", "k_radix_scatter"]}.get(kernel, [kernel])
    tot = sum(per[k] * d["launches"][k] for k in names if k in per)
    cnt = sum(d["launches"][k] for k in names if k in per)
    return (tot / cnt if cnt else None), d.get("source")


def b200_alg_bytes(n, m):
    """B_alg(n, m) = 18 n + 28 m"""
    return 18 * n + 28 * m


def make_text(ctx, wl, n, dist, rank, keep_host):
    """the bench text in HBM (n + 64 bytes); rank 0 builds it and the other ranks receive it.  Returns (S_dev, host
    copy or None)."""
    import torch
    S = torch.empty(n + 64, dtype=torch.uint8, device="cuda")
    ctx.synth_uniform_dev(S.data_ptr(), n, wl["seed"])
    torch.cuda.synchronize()
    host = None
    if wl["kind"] == "chm13":
        if rank == 0:
            import textgen as tg
            host = S[:n].cpu().numpy()
            host = tg.chm13_like(wl["seed"], n, base=host)
            S[:n].copy_(torch.from_numpy(host))
            torch.cuda.synchronize()
        if dist is not None:
            dist.broadcast(S, src=0)
            torch.cuda.synchronize()
    elif keep_host and rank == 0:
        host = S[:n].cpu().numpy()
    return S, (host if keep_host else None)


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
    n, k, wl = args.n, args.k, args.wl
    ctx = kiss_b200.Context(local)
    lib_stream = torch.cuda.ExternalStream(kiss_b200.lib().kiss_b200_stream(ctx.handle), device=local)
    want_host = rank == 0 and not args.no_e2e
    S, host_copy = make_text(ctx, wl, n, dist, rank, want_host)
    SA = torch.empty(n + 1, dtype=torch.int32, device="cuda")

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
    prof_bytes = ctx.profile_alg_bytes()
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
        "config": {"workload": f"{wl['name']} {n} bp (seed {wl['seed']}), k={k}, PARALLEL_SORTING, "
                               f"{'LMS sort sharded over %d GPUs, induction on rank 0' % world if world > 1 else '1 B200'}",
                   "baseline_config": {"c2": "configs[1]", "c3": "configs[2]"}[args.workload],
                   "n": n, "m": m, "k": k, "l2": "inputs larger than L2 (text n bytes, SA 4(n+1) bytes)"},
        "clocks": clk,
        "gpu_launches": launches,
        "stages_ms": {kk: v / args.steps * 1e3 for kk, v in stage.items()},
        "waves": [last.get("waves_L"), last.get("waves_S")],
        "lms_sort_tied": last.get("lms_sort_tied"), "lms_sort_rounds": last.get("lms_sort_rounds"),
    }
    # parity, outside the timed region: the order-sensitive checksum of the device suffix array of the last step against
    # the value tests/golden/make_full_size_golden.py computed from the REFERENCE's array of the same text (rank 0 holds
    # the array at every N, so this covers the sharded path too)
    if rank == 0 and not args.no_parity:
        got = ctx.sa_checksum_dev(SA.data_ptr(), n + 1)
        gold = None
        try:
            g = json.load(open(os.path.join(ROOT, "tests", "golden", "full_size_golden.json")))
            for e in g.values():
                if (e["n"], e["seed"], e["kind"], e["k"]) == (n, wl["seed"], wl["kind"], k):
                    gold = e["sa_mixed_sum"]
        except Exception:
            pass
        line["parity"] = {"sa_mixed_sum": f"{got:016x}", "golden": gold, "ok": (gold == f"{got:016x}") if gold else None,
                          "golden_source": "tests/golden/full_size_golden.json (reference kiss1_suffix_array_dna via oracle/_ref)",
                          "sa0_is_n": int(SA[0].item()) & 0xFFFFFFFF == n}
        # the reference's own test predicate (tests/kiss.cpp:25-33) over the whole device array: every adjacent pair ordered
        # by its first k bases, and the array a permutation of 0..n — size-independent, checked at the full size every run
        v = ctx.verify_sa_dev(S.data_ptr(), n, SA.data_ptr(), k)
        line["parity"]["k_order"] = dict(v, ok=not any(v.values()))
    if "steps_ms" in last:
        line["sharded_sort_steps_ms"] = last["steps_ms"]
    if "shard_sizes" in last:
        line["shard_sizes"] = last["shard_sizes"]
        line["sharded_concat"] = last.get("concat")

    if rank == 0:
        # roofline of the dominant kernel: algorithmic bytes per launch / its mean launch duration, measured
        # live with CUDA events around every launch (kiss_b200_profile_*), vs the measured copy bandwidth
        peaks = {}
        try:
            peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
        except Exception:
            pass
        peak = float(peaks.get("hbm_gbs", 6650.0))
        n_long = m - 1
        # one entry per kernel FUNCTION: the profiler classes of k_wave_fused / k_wave_tail are one kernel (K5), the
        # six size classes of k_groups_medium are template instances of one kernel
        def function_of(name):
            if name.startswith("k_wave"):
                return "k_wave_fused"
            if name.startswith("k_groups_medium"):
                return "k_groups_medium"
            return name
        merged, alg_tot = {}, {}
        for name, (tot_ms, cnt) in prof.items():
            key = function_of(name)
            a, b = merged.get(key, (0.0, 0))
            merged[key] = (a + tot_ms, b + cnt)
            if name in prof_bytes:  # the launcher's own count (radix passes, tie-group classes)
                alg_tot[key] = alg_tot.get(key, 0) + prof_bytes[name]
        steps = args.steps
        alg = {kk: v / merged[kk][1] for kk, v in alg_tot.items()}  # algorithmic bytes per launch
        alg.setdefault("k_pack_codes", n + n // 4)
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
            key = function_of(name)
            if key != name:
                cls = {"ms_per_step": tot_ms / steps, "launches_per_step": cnt / steps}
                if name in prof_bytes:
                    cls["alg_bytes_per_step"] = prof_bytes[name] / steps
                    cls["achieved_gbs"] = prof_bytes[name] / (tot_ms * 1e-3) / 1e9
                kernels[key].setdefault("classes", {})[name] = cls
        line["kernels"] = kernels
        dom = max((kk for kk in kernels if "achieved_gbs" in kernels[kk]), key=lambda kk: kernels[kk]["ms_per_step"])
        traffic, traffic_src = ncu_traffic(args.workload, dom) if n == wl["n"] and world == 1 else (None, None)
        line["roofline"] = {"bound": "hbm", "kernel": dom, "achieved": kernels[dom]["achieved_gbs"], "peak": peak,
                            "unit": "GB/s", "frac": kernels[dom]["frac"], "traffic": traffic, "traffic_source": traffic_src,
                            "peak_source": "MEASURED_PEAKS.json hbm_gbs (measured copy)" if peaks else "fallback 6650 GB/s",
                            "share_of_step": kernels[dom]["ms_per_step"] / ms_step,
                            "launches_per_step": kernels[dom]["launches_per_step"]}
        if "lms_sort" in stage:  # the stage the north star asks to be reported at 2/4/8 GPUs as well
            sort_s = stage["lms_sort"] / args.steps
            line["lms_sort_stage"] = {"ms": sort_s * 1e3, "m": m, "value": m / sort_s / 1e9, "unit": "G LMS suffixes/s",
                                      "alg_bytes": 16 * m, "frac": 16 * m / sort_s / 1e9 / peak}
            # the stage that carries most of the step, in the same shape as `roofline` (the largest single kernel above)
            line["roofline_lms_sort_stage"] = {"bound": "hbm", "stage": "lms_sort (K3: radix passes + tie-group kernels)",
                                               "achieved": 16 * m / sort_s / 1e9, "peak": peak, "unit": "GB/s",
                                               "frac": 16 * m / sort_s / 1e9 / peak, "share_of_step": sort_s * 1e3 / ms_step,
                                               "alg_bytes": 16 * m}
        balg = b200_alg_bytes(n, m)
        dev_s = stage.get("total_device", ms * 1e-3) / args.steps
        line["roofline_whole_path"] = {"alg_bytes": balg, "achieved": balg / dev_s / 1e9, "frac": balg / dev_s / 1e9 / peak,
                                       "note": "B_alg = 18 n + 28 m (SURVEY.md 8(d))"}

    # ---- end to end: host buffers in, host suffix array out, copies inside the timed region.  N=1: the C-ABI
    # host call.  N>1: rank 0 uploads the text, every rank receives it (NCCL broadcast), the sharded sort and the
    # induction run, rank 0 reads the suffix array back.
    if not args.no_e2e:
        import numpy as np
        hS = hSA = None
        if rank == 0:
            hS = torch.empty(n, dtype=torch.uint8).pin_memory()
            hSA = torch.empty(n + 1, dtype=torch.int32).pin_memory()
            if host_copy is not None:
                hS.copy_(torch.from_numpy(host_copy))
            else:
                hS.copy_(S[:n])
            torch.cuda.synchronize()
        reps = max(1, min(args.steps, 3))
        if world == 1:
            del SA
            torch.cuda.empty_cache()
            hS_np = hS.numpy()
            hSA_np = hSA.numpy().view(np.uint32)
            # warm-up: the first host call sizes the library's device arena for the two extra buffers (text, SA) it
            # owns on this path, the second one runs after the arena was regrown to that peak — steady state from then on
            for _ in range(2):
                ctx.suffix_array_dna(hS_np, k, out=hSA_np)
            t0 = time.perf_counter()
            for _ in range(reps):
                ctx.suffix_array_dna(hS_np, k, out=hSA_np)
            torch.cuda.synchronize()
            sec = (time.perf_counter() - t0) / reps
            assert int(hSA_np[0]) == n
            line["e2e"] = {"value": (n + 1) / sec / 1e9, "unit": "Gsuffixes/s", "h2d_bytes_per_step": n,
                           "d2h_bytes_per_step": 4 * (n + 1), "ms_per_step": sec * 1e3,
                           "h2d_ms": ctx.last_timings["h2d"] * 1e3, "d2h_ms": ctx.last_timings["d2h"] * 1e3,
                           "device_stages_ms": {kk: ctx.last_timings[kk] * 1e3 for kk in
                                                ("pack", "get_lms", "lms_sort", "put_lms", "induce")}}
            # the same text through the step AFTER the path: host codes in, the reference's .fmi
            # image out (BWT + occ + sampled SA, k = 32 as FMIndex::build fixes it) — the suffix array never crosses PCIe
            del hSA, hSA_np
            nb = kiss_b200.lib().kiss_b200_fmindex_image_bytes(n, 4)
            himg = torch.empty(nb, dtype=torch.uint8).pin_memory()
            himg_np = himg.numpy()
            for _ in range(2):
                ctx.fmindex_build(hS_np, 32, sa_intv=4, out=himg_np)
            t0 = time.perf_counter()
            for _ in range(reps):
                ctx.fmindex_build(hS_np, 32, sa_intv=4, out=himg_np)
            torch.cuda.synchronize()
            fsec = (time.perf_counter() - t0) / reps
            lt = ctx.last_timings
            line["e2e_fmindex"] = {"value": (n + 1) / fsec / 1e9, "unit": "Gsuffixes/s", "ms_per_step": fsec * 1e3,
                                   "h2d_bytes_per_step": n, "d2h_bytes_per_step": int(nb), "k": 32, "sa_intv": 4,
                                   "h2d_ms": lt["h2d"] * 1e3, "suffix_array_device_ms": lt["total_device"] * 1e3,
                                   "fmindex_kernels_ms": lt["fmindex"] * 1e3, "d2h_ms": lt["d2h"] * 1e3,
                                   "what": "kiss_b200_fmindex_build_dna_u32: host codes in, the reference's .fmi image out"}
            del himg, himg_np
            hSA = None
        else:
            # N > 1: ONE host text buffer and ONE host suffix-array buffer shared by the rank processes (POSIX shared
            # memory, page-locked in every process).  Rank r uploads slice r of the text over ITS OWN PCIe link, an NCCL
            # all-gather over NVLink completes the text on every GPU, the sharded sort and the induction run, rank 0
            # scatters the suffix array over NVLink and every rank drains its slice through its own link into the shared
            # buffer: the 15.6 GB that cross PCIe are spread over N links instead of rank 0's one.
            from multiprocessing import shared_memory
            chunk = ((n + world - 1) // world + 15) // 16 * 16
            cs = (n + 1 + world - 1) // world
            off_sa = (n + 4095) // 4096 * 4096
            size = off_sa + 4 * (n + 1)
            name = f"kiss_b200_bench_{os.environ.get('MASTER_PORT', '0')}"
            if rank == 0:
                try:
                    shared_memory.SharedMemory(name=name).unlink()
                except Exception:
                    pass
                shm = shared_memory.SharedMemory(create=True, size=size, name=name)
            dist.barrier()
            if rank != 0:
                shm = shared_memory.SharedMemory(name=name)
            hb = np.ndarray((size,), np.uint8, buffer=shm.buf)
            sh_S = torch.from_numpy(hb[:n])
            sh_SA = torch.from_numpy(hb[off_sa:off_sa + 4 * (n + 1)].view(np.int32))
            if rank == 0:
                sh_S.copy_(hS)
            del hS, hSA
            hS = hSA = None
            dist.barrier()
            Sg = torch.empty(chunk * world + 64, dtype=torch.uint8, device="cuda")   # the text, gathered
            myslice = torch.empty(cs, dtype=torch.int32, device="cuda")
            SAp = torch.empty(cs * world, dtype=torch.int32, device="cuda") if rank == 0 else None
            del S, SA
            torch.cuda.empty_cache()
            lo, hi = min(n, rank * chunk), min(n, (rank + 1) * chunk)
            slo, shi = min(n + 1, rank * cs), min(n + 1, (rank + 1) * cs)
            # page-lock what this rank moves: its slice of the text and its slice of the array (whole pages)
            registered = []
            for b0, b1 in ((lo, hi), (off_sa + 4 * slo, off_sa + 4 * shi)):
                p0, p1 = b0 // 4096 * 4096, min(size, (b1 + 4095) // 4096 * 4096)
                if p1 > p0:
                    hb[p0:p1:4096] |= 0                                   # (touch: the pages exist before they are locked)
                    rc = int(torch.cuda.cudart().cudaHostRegister(hb.ctypes.data + p0, p1 - p0, 0))
                    if rc == 0:
                        registered.append(hb.ctypes.data + p0)
                    else:
                        print(f"[bench] rank {rank}: cudaHostRegister failed with code {rc}; that slice stays pageable", file=sys.stderr)

            def e2e_step(marks=None):
                def mark(name):  # (only in the one extra, untimed step that reports the phases)
                    if marks is not None:
                        torch.cuda.synchronize()
                        dist.barrier()
                        marks.append((name, time.perf_counter()))
                mark("start")
                if hi > lo:
                    Sg[lo:hi].copy_(sh_S[lo:hi], non_blocking=True)
                mark("h2d_slice")
                dist.all_gather_into_tensor(Sg[:chunk * world], Sg[rank * chunk:(rank + 1) * chunk])
                torch.cuda.synchronize()
                mark("all_gather_text")
                sharded_suffix_array_dev(ctx, Sg, n, SAp if rank == 0 else myslice, k, dist)
                mark("suffix_array")
                dist.scatter(myslice, [SAp[i * cs:(i + 1) * cs] for i in range(world)] if rank == 0 else None, src=0)
                mark("scatter_sa")
                if shi > slo:
                    sh_SA[slo:shi].copy_(myslice[:shi - slo], non_blocking=True)
                torch.cuda.synchronize()
                mark("d2h_slice")
            for _ in range(2):  # warm-up
                e2e_step()
            barrier()
            t0 = time.perf_counter()
            for _ in range(reps):
                e2e_step()
            barrier()
            tt = torch.tensor([(time.perf_counter() - t0) / reps], device="cuda", dtype=torch.float64)
            dist.all_reduce(tt, op=dist.ReduceOp.MAX)
            sec = float(tt.item())
            marks = []
            e2e_step(marks)   # one more step with a barrier after every phase: where the time goes (rank 0's clock)
            phases = {marks[i][0]: round((marks[i][1] - marks[i - 1][1]) * 1e3, 2) for i in range(1, len(marks))}
            if rank == 0:
                # the shared host buffer holds the whole array: its first entry, and 1 M random entries against the device copy
                assert (int(sh_SA[0]) & 0xFFFFFFFF) == n
                idx = torch.randint(0, n + 1, (1_000_000,))
                assert torch.equal(sh_SA[idx], SAp[idx.cuda()].cpu()), "host suffix array differs from the device copy"
                line["e2e"] = {"value": (n + 1) / sec / 1e9, "unit": "Gsuffixes/s", "h2d_bytes_per_step": n,
                               "d2h_bytes_per_step": 4 * (n + 1), "ms_per_step": sec * 1e3,
                               "pinned": len(registered) == 2, "phases_ms": phases,
                               "torch_sees_pinned": bool(sh_S.is_pinned()),
                               "note": f"one shared host buffer (page-locked slice by slice); every rank uploads 1/{world} of the text and drains "
                                       f"1/{world} of the suffix array through its own PCIe link; NCCL all-gather of the text "
                                       "and scatter of the array over NVLink in between"}
            del sh_S, sh_SA
            for ptr in registered:
                torch.cuda.cudart().cudaHostUnregister(ptr)
            del hb
            dist.barrier()
            shm.close()
            if rank == 0:
                shm.unlink()
        hS = hSA = None
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        try:
            line["cpu_baseline"] = cpu_baseline(wl, k, args.cpu_sample)
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
