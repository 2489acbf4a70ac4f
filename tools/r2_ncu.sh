#!/bin/bash
# `ncu --set full` captures of the dominant kernels of one C3 step, summarised to text ON the GPU box (the reports are
# tens of MB each and gpurun brings back at most 64 MiB): key figures (tools/ncu_one.py) and, for the tie-group team
# kernel, instructions and stall samples per source line (tools/ncu_lines.py: ncu's SASS page joined with nvdisasm's
# line table of the built library).
set -u
TAG=${1:-r2_ncu}
OUT=gpurun_out
mkdir -p $OUT /tmp/ncu
CMD="python bench.py --steps 1 --warmup 0 --no-e2e --no-cpu-baseline --no-parity"
$CMD > $OUT/${TAG}_plain.log 2>&1 || { echo "plain run failed"; exit 1; }
cap() {  # name, kernel regex, launches to skip, launches to take
  timeout 900 ncu --set full --clock-control none --import-source on -k regex:$2 -s $3 -c $4 -f -o /tmp/ncu/$1 $CMD > $OUT/${TAG}_$1.log 2>&1
  python tools/ncu_one.py /tmp/ncu/$1.ncu-rep "== $1: ncu --set full -k regex:$2 -s $3 -c $4, C3 (3.1 Gbp), one step" > $OUT/${TAG}_$1.txt 2>&1
}
cap radix_scatter k_radix_scatter_p 1 1
cap groups_tiny k_groups_tiny 0 1
cap medium2048 k_groups_medium 5 1
cap waves k_wave_fused 1 3
cap fm_unused k_pd_key2_plain 0 1 2>/dev/null
mkdir -p /tmp/cubin && (cd /tmp/cubin && cuobjdump -xelf all $PWD/../../$GRAFT_REPO_ROOT/kiss_b200/libkiss_b200.so > /dev/null 2>&1 || cuobjdump -xelf all /root/repo/kiss_b200/libkiss_b200.so > /dev/null 2>&1)
python tools/ncu_lines.py /tmp/ncu/medium2048.ncu-rep /tmp/cubin/lms_sort.sm_100a.cubin k_groups_mediumILi16ELi4ELb0ELb0 40 > $OUT/${TAG}_medium2048_lines.txt 2>&1
python tools/ncu_lines.py /tmp/ncu/groups_tiny.ncu-rep /tmp/cubin/lms_sort.sm_100a.cubin k_groups_tinyIjE 25 > $OUT/${TAG}_groups_tiny_lines.txt 2>&1
rm -f $OUT/${TAG}_fm_unused.*
ls -la $OUT/${TAG}_* /tmp/ncu
