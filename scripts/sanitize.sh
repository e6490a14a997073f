#!/bin/bash
# compute-sanitizer over the small-shape GPU parity tests (SURVEY 5 "build adds"): memcheck, racecheck, synccheck.
#   gpurun --timeout 2400 -- 'bash scripts/sanitize.sh r02'
# Logs: gpurun_out/sanitize_<tool>_<tag>.log (copy the summaries to profiles/).  The test selection keeps every kernel family
# (streaming K1-K3, tcgen05 Gram K4/K7, MUFU K5, radix sort + GEMM K6, SIMT fallbacks) at shapes a sanitizer finishes.
set -u
T=${1:-r02}
mkdir -p gpurun_out
SEL=(
  "tests/test_gpu_stream.py::test_reparam_matches_golden_with_injected_noise"
  "tests/test_gpu_stream.py::test_vanilla_loss_golden"
  "tests/test_gpu_stream.py::test_iw_golden"
  "tests/test_gpu_stream.py::test_bce_reconstruction_mode"
  "tests/test_gpu_mmd.py::test_wae_golden"
  "tests/test_gpu_mmd.py::test_mmd_vs_oracle"
  "tests/test_gpu_mmd.py::test_mmd_row_sharded_sweeps_add_up"
  "tests/test_gpu_mmd.py::test_mmd_latent_dims_beyond_128"
  "tests/test_gpu_vq.py::test_vq_golden"
  "tests/test_gpu_vq.py::test_vq_vs_oracle"
  "tests/test_gpu_swd.py::test_swae_golden"
  "tests/test_gpu_swd.py::test_swd_other_wasserstein_degrees"
  "tests/test_gpu_betatc.py::test_betatc_golden"
  "tests/test_gpu_betatc.py::test_betatc_row_shards_add_up"
  "tests/test_gpu_pairk.py"
)
for tool in memcheck racecheck synccheck; do
  log=gpurun_out/sanitize_${tool}_$T.log
  timeout ${SAN_TIMEOUT:-700} compute-sanitizer --tool $tool --target-processes all --print-limit 20 \
      python -m pytest "${SEL[@]}" -q -x -p no:cacheprovider > $log 2>&1
  echo "$tool rc=$? : $(grep -E 'ERROR SUMMARY|RACECHECK SUMMARY|passed|failed' $log | tail -3 | tr '\n' ' ')"
done
