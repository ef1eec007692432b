# round 2d: ncu full captures of the kernels added late in round 2 (every ncu command follows a plain run of the same program)
set -x
python tools/attn_long_timeline.py > gpurun_out/plain_al.log 2>&1 || exit 1
ncu --set full --clock-control none --import-source on -k regex:attn_long_kernel -s 6 -c 1 -o gpurun_out/r02d_attn_long_prefix_b64 \
    python tools/attn_long_timeline.py > gpurun_out/ncu_al.log 2>&1
ncu -i gpurun_out/r02d_attn_long_prefix_b64.ncu-rep --page raw --csv > gpurun_out/r02d_attn_long_prefix_b64_full_raw.csv 2>/dev/null
python tools/x3_gemm_probe.py > gpurun_out/plain_x3.log 2>&1 || exit 1
ncu --set full --clock-control none --import-source on -k regex:gemm_tc_seg_kernel -s 30 -c 1 -o gpurun_out/r02d_gemm_seg_f32 \
    python tools/x3_gemm_probe.py > gpurun_out/ncu_x3.log 2>&1
ncu -i gpurun_out/r02d_gemm_seg_f32.ncu-rep --page raw --csv > gpurun_out/r02d_gemm_seg_f32_full_raw.csv 2>/dev/null
python - <<'PY'
import csv
for f in ('gpurun_out/r02d_attn_long_prefix_b64_full_raw.csv', 'gpurun_out/r02d_gemm_seg_f32_full_raw.csv'):
    rows = list(csv.reader(open(f))); h, u, v = rows[0], rows[1], rows[2]
    print(f, v[h.index('Kernel Name')][:80])
    for k in ('gpu__time_duration.sum', 'launch__grid_size', 'launch__registers_per_thread', 'dram__bytes_read.sum', 'dram__bytes_write.sum',
              'sm__ops_path_tensor_op_utchmma_src_bf16_dst_fp32_sparsity_off.avg.pct_of_peak_sustained_elapsed',
              'sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active', 'smsp__issue_active.avg.pct_of_peak_sustained_active',
              'sm__throughput.avg.pct_of_peak_sustained_elapsed', 'lts__t_sector_hit_rate.pct'):
        if k in h:
            print('  ', k, v[h.index(k)], u[h.index(k)])
PY
