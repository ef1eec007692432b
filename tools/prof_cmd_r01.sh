set -x
python tools/profile_step.py 64 2 > gpurun_out/plain64.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none -s 2200 -c 2100 --csv --log-file gpurun_out/launches_b64.csv python tools/profile_step.py 64 2 > gpurun_out/ncu_l64.log 2>&1
python tools/profile_step.py 64 1 > gpurun_out/plain64b.log 2>&1 && \
ncu --set full --clock-control none --import-source on --kernel-name-base mangled -k regex:gemm_tc_kernelILi256ELi1E -s 2 -c 2 -o gpurun_out/prof_gateup_b64 python tools/profile_step.py 64 1 > gpurun_out/ncu_f64.log 2>&1
tail -3 gpurun_out/plain64.log gpurun_out/ncu_l64.log gpurun_out/ncu_f64.log
ls -la gpurun_out
