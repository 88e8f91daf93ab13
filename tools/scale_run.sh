#!/bin/bash
# 1 -> 8 GPU scaling of bench.py on one box (the driver's round-end protocol), plus the two-device engine test.
python -m pytest tests/test_gpu_engine.py -x -q -k "second_device" 2>&1 | tail -2
python bench.py --gpus 1 --steps 60 --warmup 10 --no-extra-legs 2>gpurun_out/scale_err.log | tail -1 > gpurun_out/r02_scale_n1.json
for n in 2 4 8; do
  python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $((29600 + n)) \
    bench.py --gpus $n --steps 60 --warmup 10 --no-extra-legs 2>>gpurun_out/scale_err.log | tail -1 > gpurun_out/r02_scale_n$n.json
done
tail -2 gpurun_out/scale_err.log
