# compute-sanitizer pass over the kernel unit tests and the reduced-depth e2e cases (SURVEY.md section 5 plan item):
# memcheck everywhere, racecheck (shared-memory hazards) where the kernels hand-roll mbarrier rings / named barriers /
# DSMEM reductions. Run on the GPU box through gpurun; logs land in gpurun_out/ and are copied to profiles/.
set -x
export PI0_SANITIZE=1
SAN=/usr/local/cuda/bin/compute-sanitizer
run() {  # name tool tests...
  local name=$1 tool=$2; shift 2
  timeout 900 $SAN --tool $tool --error-exitcode 99 --print-limit 20 --log-file gpurun_out/r02_sanitizer_${name}_${tool}.log \
      python -m pytest "$@" -x -q > gpurun_out/r02_sanitizer_${name}_${tool}.pytest.log 2>&1
  echo "rc=$? ($name, $tool)" | tee -a gpurun_out/r02_sanitizer_summary.txt
  tail -3 gpurun_out/r02_sanitizer_${name}_${tool}.pytest.log | tee -a gpurun_out/r02_sanitizer_summary.txt
  grep -E "ERROR SUMMARY|RACECHECK SUMMARY" gpurun_out/r02_sanitizer_${name}_${tool}.log | tee -a gpurun_out/r02_sanitizer_summary.txt
}
rm -f gpurun_out/r02_sanitizer_summary.txt
run elementwise memcheck tests/test_elementwise_gpu.py
run gemm memcheck tests/test_gemm_gpu.py
run e2e_small memcheck tests/test_e2e_gpu.py -k "small_depth or rollout_batch or two_cameras"
run denoise_loop memcheck tests/test_denoise_loop_gpu.py -k "bitwise"
run elementwise racecheck tests/test_elementwise_gpu.py
run denoise_loop racecheck tests/test_denoise_loop_gpu.py -k "bitwise and 1-"
cat gpurun_out/r02_sanitizer_summary.txt
