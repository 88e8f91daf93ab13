mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_engine.py -m gpu -x -q > gpurun_out/pytest_part.log 2>&1; echo "pytest rc=$?"; tail -8 gpurun_out/pytest_part.log
for v in 1 2; do
timeout 400 python bench.py --steps 90 --warmup 9 --no-cpu-baseline --seqs-per-gpu 0 > gpurun_out/bench11.log 2>&1; python -c "
import json; d=json.loads(open('gpurun_out/bench11.log').read().strip().splitlines()[-1]); print('run$v', d['value'], d['e2e']['value'], d['host_enqueue_ms_per_step'], d.get('bank_indexes'))" || tail -5 gpurun_out/bench11.log
done
