mkdir -p gpurun_out
timeout 900 python -m pytest tests -x -q -m gpu > gpurun_out/pytest_gpu.log 2>&1; echo "pytest exit $?"; tail -3 gpurun_out/pytest_gpu.log
timeout 600 python tools/bench_ksweep.py > gpurun_out/ksweep.jsonl 2> gpurun_out/ksweep.err; echo "ksweep exit $?"; cat gpurun_out/ksweep.jsonl | cut -c1-200
