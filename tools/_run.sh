mkdir -p gpurun_out
for v in "RMEM_FORK_ENCODER=1" "RMEM_FORK_ENCODER=0" "RMEM_FORK_ENCODER=1" "RMEM_FORK_ENCODER=0"; do
env $v timeout 400 python bench.py --steps 90 --warmup 9 --no-cpu-baseline --seqs-per-gpu 0 > gpurun_out/bench9.log 2>&1; python -c "
import json; d=json.loads(open('gpurun_out/bench9.log').read().strip().splitlines()[-1]); print('$v', d['value'], d['e2e']['value'], d['host_enqueue_ms_per_step'])"
done
