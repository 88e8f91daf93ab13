mkdir -p gpurun_out
timeout 900 python -m pytest tests -x -q -m gpu > gpurun_out/pytest_gpu.log 2>&1; echo "pytest exit $?"; tail -3 gpurun_out/pytest_gpu.log
timeout 300 python __graft_entry__.py smoke > gpurun_out/smoke.log 2>&1; echo "smoke exit $?"; tail -1 gpurun_out/smoke.log
timeout 400 python bench.py --steps 60 --warmup 6 --no-cpu-baseline > gpurun_out/bench4.log 2>&1; echo "bench exit $?"; tail -1 gpurun_out/bench4.log
RMEM_ASYNC_WRITE=0 timeout 400 python bench.py --steps 60 --warmup 6 --no-cpu-baseline > gpurun_out/bench4b.log 2>&1; echo "bench (sync write) exit $?"; tail -1 gpurun_out/bench4b.log | cut -c1-200
timeout 300 python tools/dev_trace.py > gpurun_out/trace.log 2>&1; echo "trace exit $?"; head -4 gpurun_out/trace.log | tail -1
