#!/bin/bash
# Round-2 evidence pass on one GPU box: full parity suite, smoke, bench lines (three precisions + python executor), reference
# arm, host profile, ncu launch list, ncu --set full of the top kernels, SASS excerpts.  Usage: bash tools/gpu_check_r02.sh [tag]
TAG=${1:-r02}
O=gpurun_out
mkdir -p $O
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > $O/${TAG}_smi.txt 2>&1
python -m pytest tests -m gpu -q > $O/${TAG}_pytest_gpu.log 2>&1; echo "pytest exit $?" >> $O/${TAG}_pytest_gpu.log
tail -6 $O/${TAG}_pytest_gpu.log
python -c "import __graft_entry__ as g; g.smoke()" > $O/${TAG}_smoke.log 2>&1; echo "smoke exit $?" >> $O/${TAG}_smoke.log
tail -3 $O/${TAG}_smoke.log
python bench.py --steps 20 --warmup 5 > $O/${TAG}_bench_bf16x3.json 2> $O/${TAG}_bench_bf16x3.err; echo "bench exit $?"
cat $O/${TAG}_bench_bf16x3.json
python bench.py --steps 20 --warmup 5 --precision bf16 --no-cpu-baseline > $O/${TAG}_bench_bf16.json 2> $O/${TAG}_bench_bf16.err
cut -c1-400 $O/${TAG}_bench_bf16.json
python bench.py --steps 10 --warmup 3 --precision fp32 --no-cpu-baseline > $O/${TAG}_bench_fp32.json 2> $O/${TAG}_bench_fp32.err
cut -c1-400 $O/${TAG}_bench_fp32.json
BENCH_PYTHON_EXECUTOR=1 python bench.py --steps 20 --warmup 5 --no-cpu-baseline --no-batched-variant > $O/${TAG}_bench_python_executor.json 2> /dev/null
cut -c1-400 $O/${TAG}_bench_python_executor.json
python bench.py --impl reference --steps 6 --warmup 3 > $O/${TAG}_bench_reference.json 2> $O/${TAG}_bench_reference.err
cut -c1-300 $O/${TAG}_bench_reference.json
python tools/host_profile.py > $O/${TAG}_host_profile.txt 2>&1
head -4 $O/${TAG}_host_profile.txt
python tools/bench_fused.py bf16x3 > $O/${TAG}_fused_bf16x3.txt 2>&1; python tools/bench_fused.py bf16 > $O/${TAG}_fused_bf16.txt 2>&1
python bench.py --steps 1 --warmup 3 --no-cpu-baseline --no-batched-variant --no-train-variant > $O/${TAG}_plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 4000 --csv --log-file $O/${TAG}_launches.csv \
    python bench.py --steps 1 --warmup 3 --no-cpu-baseline --no-batched-variant --no-train-variant > $O/${TAG}_ncu.log 2>&1
echo "ncu exit $?"
python tools/summarize_launches.py $O/${TAG}_launches.csv > $O/${TAG}_launches_summary.txt 2>&1
gzip -f $O/${TAG}_launches.csv
# EVERY linear_planes launch of one forward (73: the third forward of tools/ncu_forward.py), so that the mean DRAM bytes per
# launch (bench.py's roofline.traffic) is over the same population as the algorithmic bytes per launch it is compared with
python tools/ncu_forward.py > $O/${TAG}_ncu_forward_plain.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:linear_planes_kernel -s 146 -c 73 -f -o $O/${TAG}_linear_planes \
    python tools/ncu_forward.py > $O/${TAG}_ncu_full.log 2>&1
ncu -i $O/${TAG}_linear_planes.ncu-rep --page raw --csv > $O/${TAG}_linear_planes_raw.csv 2>/dev/null
python tools/timeline.py bf16x3 > $O/${TAG}_timeline.txt 2>&1
python tools/bench_simulator.py > $O/${TAG}_simulator.txt 2>&1
python tools/bench_linear2.py bf16x3 > $O/${TAG}_linear2.txt 2>&1
ncu --set full --clock-control none --import-source on -k regex:edge_chain_kernel -s 2 -c 2 -f -o $O/${TAG}_edge_chain \
    python bench.py --steps 1 --warmup 3 --no-cpu-baseline --no-batched-variant --no-train-variant > $O/${TAG}_ncu_full2.log 2>&1
ncu -i $O/${TAG}_edge_chain.ncu-rep --page raw --csv > $O/${TAG}_edge_chain_raw.csv 2>/dev/null
rm -f $O/${TAG}_linear_planes.ncu-rep $O/${TAG}_edge_chain.ncu-rep
echo done
