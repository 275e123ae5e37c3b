#!/bin/bash
# Round 2, last part (tag r02d): the slab schedule of the greedy patch loop (dnp_slab.cu).  Plain runs first (test
# suite, bench line, probe); every ncu pass runs only after the same command exited 0 without ncu.
set -u
OUT=gpurun_out
mkdir -p $OUT
NCU="ncu --clock-control none"
( time python -m pytest tests -x -q -m gpu ) > $OUT/r02d_gpu_tests.log 2>&1; echo "gpu tests rc=$?"; tail -3 $OUT/r02d_gpu_tests.log
python bench.py > $OUT/r02d_bench.json 2> $OUT/r02d_bench.err; echo "bench rc=$?"
python tools/kernel_zoo.py slab > $OUT/r02d_kernel_zoo.jsonl 2> $OUT/r02d_kernel_zoo.err || { echo "zoo failed"; exit 1; }
cat $OUT/r02d_kernel_zoo.jsonl
$NCU --metrics gpu__time_duration.sum -c 4200 --csv --log-file $OUT/r02d_launches_slab.csv python tools/kernel_zoo.py slab > $OUT/ncu_r02d_launches.log 2>&1
echo "launch list rc=$?"
cap() {  # name  kernel-regex  skip
    $NCU --set full --import-source on -k regex:$2 -s $3 -c 1 -o $OUT/r02d_$1 -f python tools/kernel_zoo.py slab > $OUT/ncu_r02d_$1.log 2>&1
    echo "capture $1 rc=$?"
}
cap slab_field "upairs_tile_kernel" 1500
cap slab_inter slab_interactions_kernel 1
cap slab_greedy slab_greedy_kernel 1
cap slab_accum slab_accumulate_kernel 1
ls -la $OUT | grep r02d_
