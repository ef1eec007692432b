# Round-2 profiling pass, run on the GPU box through gpurun: every ncu command follows a plain run of the same program that
# exited 0. Outputs land in gpurun_out/; the summaries are copied to profiles/ (r02_*).
set -x
for B in 64 1; do
  python tools/profile_step.py $B 2 > gpurun_out/plain$B.log 2>&1 || exit 1
  N=$(grep launches_per_pass gpurun_out/plain$B.log | awk '{print $2}')
  ncu --metrics gpu__time_duration.sum --clock-control none --kernel-name-base demangled -k regex:pi0 -s $N -c $N --csv \
      --log-file gpurun_out/r02_launches_b$B.csv python tools/profile_step.py $B 2 > gpurun_out/ncu_l$B.log 2>&1
done
# full captures (--set full, one launch each): the dominant kernel and the three attention kernels that had no capture
cap() {  # name regex batch skip
  ncu --set full --clock-control none --import-source on --kernel-name-base mangled -k regex:$2 -s $4 -c 1 \
      -o gpurun_out/r02_$1 python tools/profile_step.py $3 1 > gpurun_out/ncu_$1.log 2>&1
  ncu -i gpurun_out/r02_$1.ncu-rep --page raw --csv > gpurun_out/r02_$1_full_raw.csv 2>/dev/null
}
cap gateup_pair_b64 gemm_tc_pair_kernelILi256ELi1E 64 2
cap prefill_attn_b64 prefill_attn_kernel 64 2
cap batch_attn_b64 batch_attn_kernel 64 20
cap siglip_attn_b64 siglip_attn_kernel 64 2
cap denoise_layers_b1 denoise_layers_kernel 1 1
ls -la gpurun_out | tail -20
