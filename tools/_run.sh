mkdir -p gpurun_out
timeout 200 python bench.py --kernel-only > gpurun_out/plain_k.log 2>&1 &&
timeout 600 ncu --set full --clock-control none --import-source on -k regex:'local_attn_kernel' -s 3 -c 2 -o gpurun_out/prof_local_fused python bench.py --kernel-only > gpurun_out/ncu_full3.log 2>&1; echo "ncu full local exit $?"
