#!/usr/bin/env bash
set -e
mkdir -p gpurun_out
python tools/step_profile.py --batch 256 > gpurun_out/ncu_plain.log 2>&1
for k in k_bn_pool_fwd k_bn_pool_bwd_reduce k_bn_pool_bwd_apply k_bn_bwd_apply; do
  ncu --set full --clock-control none --import-source on --profile-from-start off --kernel-name-base function \
      -k regex:^$k\$ -c 1 -f -o gpurun_out/prof_$k python tools/step_profile.py --batch 256 > gpurun_out/ncu_$k.log 2>&1 || true
done
