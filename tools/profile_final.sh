#!/bin/bash
# End-of-round evidence at the final commit: GPU test suite, smoke, bench line, launch list of the bench command and one
# full capture of its dominant kernel (each ncu pass only after the same command exited 0 without ncu).
set -u
OUT=gpurun_out
mkdir -p $OUT
NCU="ncu --clock-control none"
( time python -m pytest tests -x -q -m gpu ) > $OUT/r02e_gpu_tests.log 2>&1; echo "gpu tests rc=$?"; tail -4 $OUT/r02e_gpu_tests.log
python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -2
python bench.py > $OUT/r02e_bench.json 2> $OUT/r02e_bench.err; echo "bench rc=$?"
python bench.py --steps 1 --warmup 3 --no-cpu-baseline --no-extras --no-strong > $OUT/r02e_bench_short.json 2> $OUT/r02e_bench_short.err; echo "short bench rc=$?"
$NCU --metrics gpu__time_duration.sum -c 6000 --csv --log-file $OUT/r02e_launches_bench_raw.csv python bench.py --steps 1 --warmup 3 --no-cpu-baseline --no-extras --no-strong > $OUT/ncu_r02e_launches.log 2>&1
echo "launch list rc=$?"
$NCU --set full --import-source on -k regex:upairs_tile_kernel -s 1700 -c 1 -o $OUT/r02e_field_tile -f python bench.py --steps 1 --warmup 3 --no-cpu-baseline --no-extras --no-strong > $OUT/ncu_r02e_field_tile.log 2>&1
echo "capture rc=$?"
python tools/slab_probe.py > $OUT/r02e_slab_probe.jsonl 2>&1; echo "probe rc=$?"; tail -3 $OUT/r02e_slab_probe.jsonl | cut -c1-300
ls -la $OUT | grep r02e_
