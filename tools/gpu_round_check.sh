#!/bin/bash
# Whole round-end evidence pass on one GPU box: tools/gpu_check.sh + the training line, the per-entry-point tables of the
# batched forward and of the training step, the other BASELINE configs and the host profile.
# Usage (from the repo root, under gpurun):  bash tools/gpu_round_check.sh [tag]
TAG=${1:-r01}
O=gpurun_out
bash tools/gpu_check.sh $TAG
python tools/profile_forward.py bf16x3 6 > $O/${TAG}_entrypoints_b6.txt 2>&1
python bench.py --mode train --steps 10 --warmup 3 > $O/${TAG}_bench_train_bf16x3.json 2> $O/${TAG}_bench_train_bf16x3.err
cat $O/${TAG}_bench_train_bf16x3.json
python tools/profile_train.py bf16x3 shapenet-seg 16 > $O/${TAG}_train_entrypoints_bf16x3.txt 2>&1
python tools/bench_configs.py > $O/${TAG}_configs.jsonl 2> $O/${TAG}_configs.err
cat $O/${TAG}_configs.jsonl
python tools/host_profile.py > $O/${TAG}_host_profile.txt 2>&1
head -3 $O/${TAG}_host_profile.txt
