set -u
OUT=gpurun_out; mkdir -p $OUT
for i in 1 2; do
python bench.py --workload c2 --steps 3 --warmup 3 --no-cpu-baseline > $OUT/r1_e2e_c2_$i.json 2> $OUT/r1_e2e_c2_$i.err
python -c "
import json;d=json.loads(open('$OUT/r1_e2e_c2_$i.json').read().strip().splitlines()[-1]);print(d['ms_per_step'], d['e2e'])"
done
python tools/e2e_timing.py 2>&1 | tail -15
