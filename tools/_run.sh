mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_attention.py tests/test_gpu_engine.py -x -q -m gpu -k "local" > gpurun_out/pytest_local.log 2>&1; echo "pytest local exit $?"; tail -3 gpurun_out/pytest_local.log | cut -c1-300
timeout 200 python bench.py --kernel-only > gpurun_out/plain_k.log 2>&1; tail -1 gpurun_out/plain_k.log | python -c "
import json,sys; d=json.loads(sys.stdin.read()); print(d['roofline_local'])"
