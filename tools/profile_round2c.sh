#!/bin/bash
# Re-capture (tag r02c) of the kernels whose inner loops changed after the r02 captures: the one-launch shares kernel
# (fast/safe field fold), the multi-launch shares kernel, the lattice kernel (no clamp) and the shared-memory kernel.
# Every ncu pass runs only after the same command exited 0 without ncu.
set -u
OUT=gpurun_out
mkdir -p $OUT
NCU="ncu --clock-control none"
for w in fused shares lattice smem; do
    python tools/kernel_zoo.py $w >> $OUT/r02c_kernel_zoo.jsonl 2>> $OUT/r02c_kernel_zoo.err || { echo "$w failed"; exit 1; }
done
cat $OUT/r02c_kernel_zoo.jsonl
cap() {  # name  kernel-regex  skip  workload
    $NCU --set full --import-source on -k regex:$2 -s $3 -c 1 -o $OUT/r02c_$1 -f python tools/kernel_zoo.py $4 > $OUT/ncu_r02c_$1.log 2>&1
    echo "capture $1 rc=$?"
}
cap fused "upairs_kernel" 1 fused
cap shares "upairs_kernel" 8 shares
cap lattice lattice_sum_kernel 1 lattice
cap smem "pairs_kernel" 1 smem
ls -la $OUT | grep r02c_
