python -m pytest tests/test_gpu_train.py -q -s 2>&1 > gpurun_out/train4.log; grep -E "^E  |frozen-BN|batch-stat|golden|passed|failed" gpurun_out/train4.log | head -60
python tools/profile_train.py bf16x3 shapenet-seg 16 > gpurun_out/prof_train_seg.txt 2>&1; head -45 gpurun_out/prof_train_seg.txt
python bench.py --mode train --steps 5 --warmup 3 > gpurun_out/bench_train.json 2> gpurun_out/bench_train.err; cat gpurun_out/bench_train.json; tail -5 gpurun_out/bench_train.err
