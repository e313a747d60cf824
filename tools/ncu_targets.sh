#!/usr/bin/env bash
# ncu --set full captures of the top kernels (1 GPU, after the plain run exited 0).  Usage: tools/ncu_targets.sh
set -e
mkdir -p gpurun_out
python tools/step_profile.py --batch 256 > gpurun_out/ncu_plain.log 2>&1
for k in k_conv_igemm k_conv_wgrad k_bn_bwd_reduce; do
  ncu --set full --clock-control none --import-source on --profile-from-start off --kernel-name-base function \
      -k regex:^$k\$ -s 2 -c 3 -f -o gpurun_out/prof_$k python tools/step_profile.py --batch 256 > gpurun_out/ncu_$k.log 2>&1 || true
done
