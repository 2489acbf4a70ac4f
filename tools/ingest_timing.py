#!/usr/bin/env python
"""Ingest step at scale: a synthetic FASTA file (uniform text, 60-column lines, one record per
100 Mbp) -> codes.  Device kernels (CUDA events, file bytes resident in HBM), the same through the host call, and the
reference's own read_sequence (oracle/_ref, serial getline) on the box's host as the baseline.  One JSON line.
    python tools/ingest_timing.py [bases=1000000000]"""
import json
import os
import sys
import time

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import kiss_b200  # noqa: E402
import oracle_lib as ol  # noqa: E402


def make_fasta(ctx, n, width=60, record=100_000_000):
    d = torch.empty(n + 64, dtype=torch.uint8, device="cuda")
    ctx.synth_uniform_dev(d.data_ptr(), n, 2)
    S = d[:n].cpu().numpy()
    del d
    parts = []
    lut = np.frombuffer(b"ACGT", np.uint8)
    for r, lo in enumerate(range(0, n, record)):
        seg = lut[S[lo:lo + record]]
        full = seg.size // width * width
        body = np.concatenate([seg[:full].reshape(-1, width), np.full((full // width, 1), 10, np.uint8)], axis=1).ravel()
        parts += [np.frombuffer(b">chr%d synthetic\n" % (r + 1), np.uint8), body]
        if full < seg.size:
            parts += [seg[full:], np.full(1, 10, np.uint8)]
    return np.concatenate(parts), S


def main():
    n = int(sys.argv[1]) if len(sys.argv) > 1 else 1_000_000_000
    ctx = kiss_b200.Context(0)
    data, S = make_fasta(ctx, n)
    nbytes = data.size
    raw = torch.from_numpy(data).cuda()
    codes = torch.empty(nbytes + 16, dtype=torch.uint8, device="cuda")
    stream = torch.cuda.ExternalStream(kiss_b200.lib().kiss_b200_stream(ctx.handle))
    torch.cuda.synchronize()
    for _ in range(2):
        got_n = ctx.ingest_dev(raw.data_ptr(), nbytes, True, codes.data_ptr())
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    reps = 5
    e0.record(stream)
    for _ in range(reps):
        ctx.ingest_dev(raw.data_ptr(), nbytes, True, codes.data_ptr())
    e1.record(stream)
    torch.cuda.synchronize()
    dev_ms = e0.elapsed_time(e1) / reps
    same = bool(got_n == n and torch.equal(codes[:n], torch.from_numpy(S).cuda()))
    out = {"file_bytes": int(nbytes), "n": int(got_n), "device_ms": dev_ms, "codes_equal_text": same,
           "device_gbs_algorithmic": (2 * nbytes + n) / dev_ms / 1e6,
           "note": "algorithmic bytes = 2 reads of the file + n written"}
    if ol.have_ref() and hasattr(ol.ref(), "ref_read_sequence"):
        sample = data[: min(nbytes, 200_000_000)]
        cut = int(np.nonzero(sample == 10)[0][-1]) + 1
        t0 = time.perf_counter()
        ref_codes = ol.ref_read_sequence(sample[:cut].tobytes())
        sec = time.perf_counter() - t0
        out["reference_read_sequence"] = {"bytes": cut, "seconds": sec, "gbs": cut / sec / 1e9, "threads": 1,
                                          "n": int(ref_codes.size)}
        out["speedup_vs_reference_ingest"] = (cut / sec) / (nbytes / (dev_ms * 1e-3))  # < 1 ... inverted below
        out["speedup_vs_reference_ingest"] = (nbytes / (dev_ms * 1e-3)) / (cut / sec)
    print(json.dumps(out), flush=True)
    ctx.close()


if __name__ == "__main__":
    main()
