mkdir -p gpurun_out
timeout 400 python tools/dev_trace.py --prefetch > gpurun_out/trace_pf.txt 2>&1; echo rc=$?; tail -3 gpurun_out/trace_pf.txt
