#!/bin/bash
# One round of evidence on the GPU box (run through gpurun): tests, bench (both arms), ncu launch list of one
# step, one `--set full` capture of the dominant kernels.  Outputs under gpurun_out/; summaries are copied to
# profiles/ by hand (tools/ncu_summary.py, tools/ncu_kernels.py, tools/ncu_dram_per_kernel.py).
set -u
TAG=${1:-r1_final}
OUT=gpurun_out
mkdir -p $OUT
python -m pytest tests -m gpu -x -q 2>&1 | tail -3 | tee $OUT/${TAG}_pytest.txt
python bench.py --impl reference --steps 2 --warmup 1 > $OUT/${TAG}_bench_reference.json 2> $OUT/${TAG}_bench_reference.err
tail -c 400 $OUT/${TAG}_bench_reference.json
python bench.py > $OUT/${TAG}_bench.json 2> $OUT/${TAG}_bench.err
tail -c 300 $OUT/${TAG}_bench.json
# launch list of the timed step (warm-up 3 steps x 96 launches + the text generator = 289 launches skipped)
ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -s 289 -c 96 --csv \
    --log-file $OUT/${TAG}_launches.csv python bench.py --steps 1 --warmup 3 --no-e2e --no-cpu-baseline > $OUT/${TAG}_ncu_launches.log 2>&1
# full captures: radix scatter (2nd pass), an LMS wave and the largest streaming wave
ncu --set full --clock-control none --import-source on -k regex:k_radix_scatter_p -s 1 -c 1 -f -o $OUT/${TAG}_radix_scatter \
    python tools/validate_large.py c2 --noref > $OUT/${TAG}_ncu_radix.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:k_wave_fused -s 4 -c 2 -f -o $OUT/${TAG}_waves \
    python tools/validate_large.py c2 --noref > $OUT/${TAG}_ncu_waves.log 2>&1
ls -la $OUT/${TAG}_*
