# round 2, final state: ncu launch lists of one eager pi0_infer at batch 64 and batch 1 (after a plain run of the same program)
set -x
for B in 64 1; do
  python tools/profile_step.py $B 2 > gpurun_out/plain_e$B.log 2>&1 || exit 1
  N=$(grep launches_per_pass gpurun_out/plain_e$B.log | awk '{print $2}')
  ncu --metrics gpu__time_duration.sum --clock-control none --kernel-name-base demangled -k regex:pi0 -s $N -c $N --csv \
      --log-file gpurun_out/r02e_launches_b$B.csv python tools/profile_step.py $B 2 > gpurun_out/ncu_e$B.log 2>&1
done
