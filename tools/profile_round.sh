#!/bin/bash
# Round profile on the B200 box: bench (plain), then the ncu launch list and one full capture of the
# dominant kernel.  Every ncu pass runs only after the same command exited 0 without ncu.
set -u
OUT=gpurun_out
mkdir -p $OUT
python bench.py --steps 3 --warmup 3 > $OUT/bench.json 2> $OUT/bench.err || { echo "bench failed"; tail -20 $OUT/bench.err; exit 1; }
cat $OUT/bench.json
python bench.py --steps 1 --warmup 3 --no-cpu-baseline --no-extras > $OUT/bench_short.json 2>&1 || exit 1
ncu --metrics gpu__time_duration.sum --clock-control none -c 6000 --csv --log-file $OUT/launches.csv \
    python bench.py --steps 1 --warmup 3 --no-cpu-baseline --no-extras > $OUT/ncu_launches.log 2>&1
echo "launch list rc=$?"
ncu --set full --clock-control none --import-source on -k regex:cpairs_kernel -s 1700 -c 1 -o $OUT/cpairs_full -f \
    python bench.py --steps 1 --warmup 3 --no-cpu-baseline --no-extras > $OUT/ncu_full.log 2>&1
echo "full capture rc=$?"
ls -la $OUT
# ---- this round's additional kernels: flip solve (exhaustive walk at 36 patches), text IO (parse / format at 1M points)
python tools/flip_probe.py > $OUT/flip_probe.json 2> $OUT/flip_probe.err || { echo "flip probe failed"; tail -5 $OUT/flip_probe.err; }
ncu --set full --clock-control none --import-source on -k regex:flip_exhaustive_kernel -s 18 -c 1 -o $OUT/flip_full -f \
    python tools/flip_probe.py > $OUT/ncu_flip.log 2>&1
echo "flip capture rc=$?"
python tools/io_gpu_probe.py > $OUT/io_probe.json 2> $OUT/io_probe.err || { echo "io probe failed"; tail -5 $OUT/io_probe.err; }
ncu --set full --clock-control none --import-source on -k regex:parse_lines_kernel -s 2 -c 1 -o $OUT/parse_full -f \
    python tools/io_gpu_probe.py > $OUT/ncu_parse.log 2>&1
echo "parse capture rc=$?"
ncu --set full --clock-control none --import-source on -k regex:format_values_kernel -s 2 -c 1 -o $OUT/format_full -f \
    python tools/io_gpu_probe.py > $OUT/ncu_format.log 2>&1
echo "format capture rc=$?"
ls -la $OUT
