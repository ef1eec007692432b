# round 2c: ncu evidence for the tcgen05 SigLIP attention (launch list of one batch-64 step + one full capture)
set -x
python tools/profile_step.py 64 2 > gpurun_out/plain64c.log 2>&1 || exit 1
N=$(grep launches_per_pass gpurun_out/plain64c.log | awk '{print $2}')
ncu --metrics gpu__time_duration.sum --clock-control none --kernel-name-base demangled -k regex:pi0 -s $N -c $N --csv \
    --log-file gpurun_out/r02c_launches_b64.csv python tools/profile_step.py 64 2 > gpurun_out/ncu_launch_c.log 2>&1
[ -n "$LAUNCHES_ONLY" ] && exit 0
ncu --set full --clock-control none --import-source on -k regex:siglip_attn_tc_kernel -s 3 -c 1 \
    -o gpurun_out/r02c_siglip_attn_tc_b64 python tools/profile_step.py 64 1 > gpurun_out/ncu_sa.log 2>&1
ncu -i gpurun_out/r02c_siglip_attn_tc_b64.ncu-rep --page raw --csv > gpurun_out/r02c_siglip_attn_tc_b64_full_raw.csv 2>/dev/null
python - <<'PY'
import csv
rows=list(csv.reader(open('gpurun_out/r02c_siglip_attn_tc_b64_full_raw.csv'))); h,u,v=rows[0],rows[1],rows[2]
for k in ('gpu__time_duration.sum','dram__bytes_read.sum','dram__bytes_write.sum','lts__t_sector_hit_rate.pct','launch__registers_per_thread',
          'sm__ops_path_tensor_op_utchmma_src_bf16_dst_fp32_sparsity_off.avg.pct_of_peak_sustained_elapsed','sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active',
          'sm__throughput.avg.pct_of_peak_sustained_elapsed','smsp__issue_active.avg.pct_of_peak_sustained_active','sm__warps_active.avg.pct_of_peak_sustained_active'):
    if k in h: print(k, v[h.index(k)], u[h.index(k)])
PY
