mkdir -p gpurun_out
timeout 900 ncu --profile-from-start off -k regex:gemm_bf16_tn_kernel -c 23 --set full --clock-control none --import-source on -f -o gpurun_out/r01_gemm python bench.py --steps 5 --warmup 3 --profile-steps 1 --no-cpu-baseline --seqs-per-gpu 0 > gpurun_out/ncu_gemm.log 2>&1; echo "ncu rc=$?"; tail -2 gpurun_out/ncu_gemm.log | cut -c1-200; ls -la gpurun_out/*.ncu-rep
