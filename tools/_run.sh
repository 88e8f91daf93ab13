timeout 900 python -m pytest tests -x -q -m gpu > gpurun_out/pytest_gpu.log 2>&1; echo "pytest exit $?"; tail -5 gpurun_out/pytest_gpu.log
timeout 300 python bench.py --steps 40 --warmup 6 --no-cpu-baseline > gpurun_out/bench5.log 2>&1; echo "bench exit $?"; tail -1 gpurun_out/bench5.log | python -c "
import json,sys
d=json.loads(sys.stdin.read()); print(d['value'], d['ms_per_step'], d['e2e']['value'], d['roofline']['frac'], d['roofline_local']['frac'], d['roofline_local']['us'])"
