#!/bin/bash
# One GPU-box pass: parity tests, bench lines (three precisions), per-entry-point timing, ncu launch list.
# Usage (from the repo root, under gpurun):  bash tools/gpu_check.sh [tag]
TAG=${1:-r01}
O=gpurun_out
mkdir -p $O
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > $O/${TAG}_smi.txt 2>&1
python -m pytest tests -m gpu -x -q > $O/${TAG}_pytest_gpu.log 2>&1; echo "pytest exit $?" >> $O/${TAG}_pytest_gpu.log
tail -5 $O/${TAG}_pytest_gpu.log
python -c "import __graft_entry__ as g; g.smoke()" > $O/${TAG}_smoke.log 2>&1; echo "smoke exit $?" >> $O/${TAG}_smoke.log
tail -2 $O/${TAG}_smoke.log
python bench.py --steps 20 --warmup 5 > $O/${TAG}_bench_bf16x3.json 2> $O/${TAG}_bench_bf16x3.err; echo "bench exit $?"
cat $O/${TAG}_bench_bf16x3.json
python bench.py --steps 20 --warmup 5 --precision bf16 --no-cpu-baseline > $O/${TAG}_bench_bf16.json 2> $O/${TAG}_bench_bf16.err
cat $O/${TAG}_bench_bf16.json
python bench.py --steps 10 --warmup 3 --precision fp32 --no-cpu-baseline > $O/${TAG}_bench_fp32.json 2> $O/${TAG}_bench_fp32.err
cat $O/${TAG}_bench_fp32.json
python tools/profile_forward.py bf16x3 > $O/${TAG}_entrypoints_bf16x3.txt 2>&1
python tools/profile_forward.py bf16 > $O/${TAG}_entrypoints_bf16.txt 2>&1
head -40 $O/${TAG}_entrypoints_bf16x3.txt
python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-batched-variant > $O/${TAG}_plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 6000 --csv --log-file $O/${TAG}_launches.csv \
    python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-batched-variant > $O/${TAG}_ncu.log 2>&1
echo "ncu exit $?"
python tools/summarize_launches.py $O/${TAG}_launches.csv > $O/${TAG}_launches_summary.txt 2>&1
gzip -f $O/${TAG}_launches.csv
# one --set full capture of the dominant kernel (a few launches from the middle of a forward)
ncu --set full --clock-control none --import-source on -k regex:linear_planes_kernel -s 60 -c 6 -f -o $O/${TAG}_linear_planes \
    python bench.py --steps 1 --warmup 3 --no-cpu-baseline --no-batched-variant > $O/${TAG}_ncu_full.log 2>&1
echo "ncu full exit $?"
ncu -i $O/${TAG}_linear_planes.ncu-rep --page raw --csv > $O/${TAG}_linear_planes_raw.csv 2>/dev/null
