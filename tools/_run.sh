mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_engine.py -m gpu -x -q > gpurun_out/pytest_part.log 2>&1; echo "pytest rc=$?"; tail -8 gpurun_out/pytest_part.log
for flags in "--no-prefetch" ""; do
timeout 400 python bench.py --steps 60 --warmup 9 --no-cpu-baseline --seqs-per-gpu 0 $flags > gpurun_out/bench11.log 2>&1; python -c "
import json; d=json.loads(open('gpurun_out/bench11.log').read().strip().splitlines()[-1]); print('flags=[$flags]', round(d['value'],1), round(d['e2e']['value'],1), d.get('bank_indexes'))" || tail -5 gpurun_out/bench11.log
done
