#!/bin/bash
# Round-2 ncu evidence: the launch list of ten steady-state frames and full captures of the attention kernels.
set -x
B="python bench.py --steps 12 --warmup 4 --profile-steps 10 --no-extra-legs --no-cpu-baseline"
$B > gpurun_out/r02_profile_plain.json 2> gpurun_out/r02_profile_plain.err || exit 1
ncu --profile-from-start off --metrics gpu__time_duration.sum --clock-control none --csv \
    --log-file gpurun_out/r02_launches_bench_steps10.csv $B > gpurun_out/r02_profile_ncu.json 2> gpurun_out/r02_profile_ncu.err
python bench.py --kernel-only > gpurun_out/r02_kernel_only.json 2> gpurun_out/r02_kernel_only.err || exit 1
for k in lt_attn_sk_kernel lt_attn_sk_combine_kernel lt_attn_partial_kernel local_scores_kernel local_aggregate_kernel; do
  ncu --set full --clock-control none --import-source on -k regex:$k --launch-skip 6 -c 2 -f \
      -o gpurun_out/r02_full_$k python bench.py --kernel-only > /dev/null 2> gpurun_out/r02_full_$k.err
done
ls -la gpurun_out/*.ncu-rep
