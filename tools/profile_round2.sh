#!/bin/bash
# Round-2 profile on the B200 box (one gpurun call).  Every ncu pass runs only after the same command exited 0
# without ncu.  Raw outputs go to gpurun_out/; tools/summarize_ncu.py turns them into the tracked files of profiles/.
set -u
OUT=gpurun_out
mkdir -p $OUT
NCU="ncu --clock-control none"
# ---- plain runs first
python bench.py --steps 1 --warmup 3 --no-cpu-baseline --no-extras --no-strong > $OUT/r02_bench_short.json 2> $OUT/r02_bench_short.err || { echo "bench failed"; tail -5 $OUT/r02_bench_short.err; exit 1; }
python tools/kernel_zoo.py all > $OUT/r02_kernel_zoo.jsonl 2> $OUT/r02_kernel_zoo.err || { echo "zoo failed"; tail -5 $OUT/r02_kernel_zoo.err; exit 1; }
cat $OUT/r02_kernel_zoo.jsonl
# ---- launch list of the bench step (cold-cache, serialised launches: shares, not absolute times)
$NCU --metrics gpu__time_duration.sum -c 6000 --csv --log-file $OUT/r02_launches.csv \
    python bench.py --steps 1 --warmup 3 --no-cpu-baseline --no-extras --no-strong > $OUT/ncu_launches.log 2>&1
echo "launch list rc=$?"
# ---- DRAM bytes of every launch of ONE field call WITHOUT the cache flush ncu does by default: what the step really
#      moves (targets and partial rows stay in the 126 MB L2 from launch to launch)
$NCU --cache-control none --metrics dram__bytes_read.sum,dram__bytes_write.sum,gpu__time_duration.sum -k regex:upairs_tile_kernel -s 1171 -c 1171 --csv \
    --log-file $OUT/r02_dram_warm.csv python tools/kernel_zoo.py field > $OUT/ncu_dram.log 2>&1
echo "warm dram rc=$?"
# ---- one full capture per kernel
cap() {  # name  kernel-regex  skip  workload
    $NCU --set full --import-source on -k regex:$2 -s $3 -c 1 -o $OUT/r02_$1 -f python tools/kernel_zoo.py $4 > $OUT/ncu_$1.log 2>&1
    echo "capture $1 rc=$?"
}
cap field_tile upairs_tile_kernel 1700 field
cap potential_tile upairs_tile_kernel 1200 potential
cap shares "upairs_kernel" 8 shares
cap fused "upairs_kernel" 1 fused
cap lattice lattice_sum_kernel 1 lattice
cap points_cluster points_propagate_kernel 1 points
cap points_grid points_propagate_grid_kernel 0 points_grid
cap edges edge_interactions_kernel 1 edges
cap patch_inter patch_interactions_kernel 1 patch_inter
cap field_sign field_sign_kernel 1 sign
cap xie_chain xie_chain_kernel 1 xie
cap cell_key cell_key_kernel 1 partition
ls -la $OUT | grep r02_
