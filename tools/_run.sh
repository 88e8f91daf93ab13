mkdir -p gpurun_out
timeout 900 python -m pytest tests -x -q -m gpu > gpurun_out/pytest_gpu.log 2>&1; echo "pytest exit $?"; tail -3 gpurun_out/pytest_gpu.log
timeout 400 python bench.py --steps 60 --warmup 6 --no-cpu-baseline > gpurun_out/bench6.log 2>&1; echo "bench exit $?"; python -c "
import json; d=json.loads(open('gpurun_out/bench6.log').read().strip().splitlines()[-1]); print(d['value'], d['e2e']['value'], d['host_enqueue_ms_per_step'])"
