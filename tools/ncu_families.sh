#!/usr/bin/env bash
# One `ncu --set full` capture per hot kernel family of the flagship step (batch 1200, fp16), two launches each.
# Run under gpurun AFTER the same command has exited 0 without ncu (B200_PROFILING.md):
#   gpurun -- 'python tools/step_profile.py --batch 1200 > gpurun_out/plain.log 2>&1 && bash tools/ncu_families.sh'
# Every capture is summarised on the box (metric table + top stalled instructions); only the reports named in
# KEEP stay (gpurun brings back at most 64 MiB).
set -u
OUT=${OUT:-gpurun_out}
BATCH=${BATCH:-1200}
KEEP=${KEEP:-"k_conv_igemm k_conv3x3_halo"}
: > $OUT/ncu_r2_summary.md
for fam in k_conv_igemm k_conv_wgrad k_conv3x3_halo k_wgrad3x3_halo k_stem_conv k_fwd_apply k_bwd_reduce k_bwd_apply \
           k_bn_pool_fwd k_bn_pool_bwd_apply k_reduce_scatter k_sgd_allgather k_ce_fwd k_dgrad_weight_batched; do
  ncu --set full --clock-control none --import-source on --profile-from-start off -k regex:$fam -c ${COUNT:-2} \
      -o $OUT/prof_r2_$fam -f python tools/step_profile.py --batch $BATCH > $OUT/ncu_r2_$fam.log 2>&1
  rc=$?
  python tools/ncu_summary.py $OUT/prof_r2_$fam.ncu-rep >> $OUT/ncu_r2_summary.md 2>>$OUT/ncu_r2_errors.txt
  python tools/ncu_hot.py $OUT/prof_r2_$fam.ncu-rep 10 > $OUT/ncu_r2_hot_$fam.txt 2>>$OUT/ncu_r2_errors.txt
  case " $KEEP " in *" $fam "*) ;; *) rm -f $OUT/prof_r2_$fam.ncu-rep ;; esac
  tail -c 400 $OUT/ncu_r2_$fam.log > $OUT/ncu_r2_$fam.tail; rm -f $OUT/ncu_r2_$fam.log
  echo "$fam exit $rc"
done
