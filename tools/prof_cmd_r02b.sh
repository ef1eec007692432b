set -x
python tools/profile_step.py 64 1 > gpurun_out/plain64b.log 2>&1 || exit 1
ncu --set full --clock-control none --import-source on --kernel-name-base mangled -k regex:gemm_tc_pair_kernelILi256ELi1E -s 2 -c 1 \
    -o gpurun_out/r02b_gateup_pair_b64 python tools/profile_step.py 64 1 > gpurun_out/ncu_gu2.log 2>&1
ncu -i gpurun_out/r02b_gateup_pair_b64.ncu-rep --page raw --csv > gpurun_out/r02b_gateup_pair_b64_full_raw.csv 2>/dev/null
python - <<'PY'
import csv
rows=list(csv.reader(open('gpurun_out/r02b_gateup_pair_b64_full_raw.csv'))); h,u,v=rows[0],rows[1],rows[2]
for k in ('gpu__time_duration.sum','dram__bytes_read.sum','dram__bytes_write.sum','lts__t_sector_hit_rate.pct','sm__ops_path_tensor_op_utchmma_src_bf16_dst_fp32_sparsity_off.avg.pct_of_peak_sustained_elapsed'):
    print(k, v[h.index(k)], u[h.index(k)])
PY
