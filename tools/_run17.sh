set -u
OUT=gpurun_out; mkdir -p $OUT
timeout 900 python tools/validate_large.py c3h --profile > $OUT/r1_c3h_validate.jsonl 2> $OUT/r1_c3h_validate.err
cut -c1-900 $OUT/r1_c3h_validate.jsonl; tail -3 $OUT/r1_c3h_validate.err
