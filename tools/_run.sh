mkdir -p gpurun_out
timeout 900 python -m pytest tests -x -q -m gpu > gpurun_out/pytest_gpu.log 2>&1; echo "pytest exit $?"; tail -3 gpurun_out/pytest_gpu.log
for m in 0 1 2; do RMEM_LT_EXP=$m timeout 200 python bench.py --kernel-only > gpurun_out/lt_exp_$m.log 2>&1; echo "exp $m: $(tail -1 gpurun_out/lt_exp_$m.log | cut -c1-260)"; done
timeout 400 python bench.py --steps 60 --warmup 6 --no-cpu-baseline > gpurun_out/bench3.log 2>&1; echo "bench exit $?"; tail -1 gpurun_out/bench3.log
