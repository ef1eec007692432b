# Profiling pass after the CTA-pair GEMM landed, run on the GPU box through gpurun: every ncu command follows a plain run
# of the same program that exited 0. Outputs land in gpurun_out/; the summaries are copied to profiles/.
set -x
for B in 64 1; do
  python tools/profile_step.py $B 2 > gpurun_out/plain$B.log 2>&1 || exit 1
  N=$(grep launches_per_pass gpurun_out/plain$B.log | awk '{print $2}')
  # launch list of the SECOND pass: only this library's kernels (namespace pi0), skip the first pass
  ncu --metrics gpu__time_duration.sum --clock-control none --kernel-name-base demangled -k regex:pi0 -s $N -c $N --csv \
      --log-file gpurun_out/r01c_launches_b$B.csv python tools/profile_step.py $B 2 > gpurun_out/ncu_l$B.log 2>&1
done
# full capture: dominant kernel of the batch-64 step (prefill gate/up GeGLU GEMM, CTA-pair kernel). (An earlier version of
# this script also captured SigLIP fc1 on the single-CTA kernel for the epilogue analysis; those GEMMs now run on the
# pair kernel with the TMA-store epilogue.)
ncu --set full --clock-control none --import-source on --kernel-name-base mangled -k regex:gemm_tc_pair_kernelILi256ELi1E -s 2 -c 1 \
    -o gpurun_out/r01c_gateup_pair_b64 python tools/profile_step.py 64 1 > gpurun_out/ncu_f64.log 2>&1
tail -2 gpurun_out/plain64.log gpurun_out/plain1.log gpurun_out/ncu_l64.log gpurun_out/ncu_f64.log
ls -la gpurun_out | tail -12
