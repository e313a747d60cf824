#!/usr/bin/env bash
# Profile evidence at HEAD (round 2, second half), summarised ON the box (gpurun brings back <= 64 MiB):
#   1. launch lists (gpu__time_duration per launch) of one eager step: fp16 at 1200 and 150 images, fp32 at 1200
#   2. one table row per launch of the fp16 / 1200 step from the SpeedOfLight + LaunchStats + MemoryWorkloadAnalysis sections
#   3. `ncu --set full --import-source on` of the kernels that changed, with the top stalled instructions of each
# Run under gpurun; every profiled command first exits 0 without ncu (B200_PROFILING.md).
set -u
OUT=${OUT:-gpurun_out}
mkdir -p $OUT
python tools/step_profile.py --batch 1200 > $OUT/r3_plain.log 2>&1 || { echo "plain run failed"; tail -5 $OUT/r3_plain.log; exit 1; }
python tools/step_profile.py --batch 1200 --mode ddp > $OUT/r3_plain32.log 2>&1 || { echo "plain fp32 run failed"; exit 1; }

# ---- 1. launch lists
for cfg in "1200 ddp_amp fp16" "150 ddp_amp fp16" "1200 ddp fp32"; do
  set -- $cfg
  ncu --profile-from-start off --metrics gpu__time_duration.sum --clock-control none --csv --log-file $OUT/r3_l.csv \
      python tools/step_profile.py --batch $1 --mode $2 > $OUT/r3_l.log 2>&1
  python tools/summarize_launches.py $OUT/r3_l.csv 60 > $OUT/launches_b$1_$3_r3.txt
  python tools/summarize_launches.py $OUT/r3_l.csv list > $OUT/launches_b$1_$3_r3_list.txt
  rm -f $OUT/r3_l.csv
  echo "launch list $cfg done"
done

# ---- 2. every launch of the fp16 step, light sections
ncu --profile-from-start off --clock-control none --section SpeedOfLight --section LaunchStats --section MemoryWorkloadAnalysis \
    --section Occupancy -f -o $OUT/r3_step python tools/step_profile.py --batch 1200 > $OUT/r3_step.log 2>&1
python tools/ncu_summary.py $OUT/r3_step.ncu-rep > $OUT/ncu_r3_step_b1200_fp16.md 2>$OUT/r3_step_sum.err
rm -f $OUT/r3_step.ncu-rep
echo "step table done: $(wc -l < $OUT/ncu_r3_step_b1200_fp16.md) rows"

# ---- 3. full captures of the changed kernels: "family skip count [mode]"
: > $OUT/ncu_r3_full.md
while read fam skip count mode; do
  [ -z "$fam" ] && continue
  ncu --set full --clock-control none --import-source on --profile-from-start off -k regex:$fam -s $skip -c $count \
      -f -o $OUT/r3_full_$fam python tools/step_profile.py --batch 1200 --mode ${mode:-ddp_amp} > $OUT/r3_full_$fam.log 2>&1
  rc=$?
  python tools/ncu_summary.py $OUT/r3_full_$fam.ncu-rep | tail -n +3 >> $OUT/ncu_r3_full.md 2>>$OUT/r3_full.err
  python tools/ncu_hot.py $OUT/r3_full_$fam.ncu-rep 12 > $OUT/ncu_r3_hot_${fam}_${mode:-ddp_amp}.txt 2>>$OUT/r3_full.err
  rm -f $OUT/r3_full_$fam.ncu-rep
  tail -c 300 $OUT/r3_full_$fam.log > $OUT/r3_full_$fam.tail; rm -f $OUT/r3_full_$fam.log
  echo "$fam ($mode) exit $rc"
done <<'EOF'
k_conv_igemm 4 2
k_conv3x3_halo 3 2
k_conv_wgrad 1 2
k_wgrad3x3_halo 0 1
k_stem_wgrad 0 1
k_bwd_reduce 16 2
k_fwd_apply 0 2
k_conv3x3_halo 0 1 ddp
k_wgrad3x3_halo_tf32 0 1 ddp
k_bn_pool_bwd_apply_321_f32 0 1 ddp
EOF
echo NCU_R3_DONE
