set -u
OUT=gpurun_out; mkdir -p $OUT
timeout 1500 python tools/validate_large.py c5k256 c5k256pd c5k128 c5k128pd c5k64 c5k64pd c5k32 c5k32pd > $OUT/r1_c5_sweep.jsonl 2> $OUT/r1_c5_sweep.err
cut -c1-400 $OUT/r1_c5_sweep.jsonl; tail -5 $OUT/r1_c5_sweep.err
