mkdir -p gpurun_out
for s in 2 3 4; do
timeout 400 python bench.py --steps 60 --warmup 6 --no-cpu-baseline --seqs-per-gpu $s > gpurun_out/bench_s$s.log 2>&1; python -c "
import json; d=json.loads(open('gpurun_out/bench_s$s.log').read().strip().splitlines()[-1]); print('S=$s', d['value'], d['multi_sequence']['value'])"
done
