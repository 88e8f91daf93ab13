mkdir -p gpurun_out
timeout 900 ncu --profile-from-start off --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/launches.csv python bench.py --steps 12 --warmup 3 --profile-steps 10 --no-cpu-baseline --seqs-per-gpu 0 > gpurun_out/ncu.log 2>&1; echo "ncu rc=$?"; tail -2 gpurun_out/ncu.log | cut -c1-200
