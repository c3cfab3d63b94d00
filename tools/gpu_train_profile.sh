#!/bin/bash
# ncu evidence for the training step: launch list (gpu__time_duration) of one bench step + --set full of the tcgen05 weight-gradient kernel
TAG=${1:-r01d}
O=gpurun_out
mkdir -p $O
python bench.py --mode train --steps 1 --warmup 3 --no-cpu-baseline > $O/${TAG}_train_plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 9000 --csv --log-file $O/${TAG}_train_launches.csv \
    python bench.py --mode train --steps 1 --warmup 3 --no-cpu-baseline > $O/${TAG}_train_ncu.log 2>&1
echo "ncu launch list exit $?"
python tools/summarize_launches.py $O/${TAG}_train_launches.csv > $O/${TAG}_train_launches_summary.txt 2>&1
gzip -f $O/${TAG}_train_launches.csv
head -30 $O/${TAG}_train_launches_summary.txt
ncu --set full --clock-control none --import-source on -k regex:wgrad_tc_kernel -s 100 -c 4 -f -o $O/${TAG}_wgrad_tc \
    python bench.py --mode train --steps 1 --warmup 3 --no-cpu-baseline > $O/${TAG}_train_ncu_full.log 2>&1
echo "ncu full exit $?"
ncu -i $O/${TAG}_wgrad_tc.ncu-rep --page raw --csv > $O/${TAG}_wgrad_tc_raw.csv 2>/dev/null
