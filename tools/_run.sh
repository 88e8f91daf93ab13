mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_engine.py -x -q -m gpu > gpurun_out/pytest_gpu.log 2>&1; echo "pytest exit $?"; tail -3 gpurun_out/pytest_gpu.log
timeout 400 python bench.py --steps 90 --warmup 9 --no-cpu-baseline --seqs-per-gpu 0 > gpurun_out/bench10.log 2>&1; echo "bench exit $?"; python -c "
import json; d=json.loads(open('gpurun_out/bench10.log').read().strip().splitlines()[-1]); print('prefetch', d['value'], d['e2e']['value'], d['host_enqueue_ms_per_step'], d['bank_indexes'])"; tail -3 gpurun_out/bench10.log | cut -c1-300
timeout 400 python bench.py --steps 90 --warmup 9 --no-cpu-baseline --seqs-per-gpu 0 --no-prefetch > gpurun_out/bench10b.log 2>&1; python -c "
import json; d=json.loads(open('gpurun_out/bench10b.log').read().strip().splitlines()[-1]); print('no prefetch', d['value'], d['e2e']['value'], d['bank_indexes'])"
