#!/bin/bash
# A/B on one box: each line = one bench.py run (C3, 5 steps) with the given environment
run() { env "$@" python bench.py --no-e2e --no-cpu-baseline --no-parity > gpurun_out/ab_tmp.json 2>/dev/null; python - "$*" <<'PY'
import json, sys
d = json.load(open("gpurun_out/ab_tmp.json"))
k = d.get("kernels", {})
def g(n):
    v = k.get(n, {})
    return round(v.get("ms_per_step", 0), 2)
t = d.get("timings", d.get("phases_ms", {}))
print(sys.argv[1], "| step", d["ms_per_step"], "| medium", g("k_groups_medium"), "tiny", g("k_groups_tiny"), "scatter", g("k_radix_scatter"), flush=True)
PY
}
for i in 1 2; do
run A=1
run KISS_B200_TEAM32=1
run KISS_B200_ONE_STREAM=1
run KISS_B200_ONE_STREAM=1 KISS_B200_TEAM32=1
done
