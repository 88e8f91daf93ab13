mkdir -p gpurun_out
timeout 200 ncu --profile-from-start off -k "regex:^[a-z0-9_]+_kernel$" -c 84 --section SpeedOfLight --section LaunchStats --section MemoryWorkloadAnalysis --clock-control none -f -o /tmp/r01_frame python bench.py --steps 5 --warmup 3 --profile-steps 1 --no-cpu-baseline --seqs-per-gpu 0 > gpurun_out/ncu_frame.log 2>&1; echo "ncu rc=$?"
ncu -i /tmp/r01_frame.ncu-rep --page raw --csv > gpurun_out/r01_frame_raw.csv 2>/dev/null; ls -la gpurun_out/r01_frame_raw.csv
