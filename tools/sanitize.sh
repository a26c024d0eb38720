#!/bin/bash
# compute-sanitizer over a trimmed selection of the -m gpu tests: the three
# hand-rolled protocols — in-launch last-arriver merge, generic-proxy scatter -> fence.proxy.async
# -> TMA read, PDL (griddepcontrol) — plus cascade, bulk norm-rope, tcgen05 prefill and the store.
#   tools/sanitize.sh memcheck|racecheck|synccheck|initcheck [extra pytest args]
# Writes gpurun_out/r02_sanitizer_<tool>.txt (sanitizer log) and ..._pytest.txt. One tool per call.
# Allocator caching is off so memcheck sees the real bounds of every tensor.
# NOTE (round 2): the GPU pool refuses compute-sanitizer (profiles/r02_sanitizer_unavailable.txt).
set -u
TOOL=${1:-memcheck}; shift || true
OUT=gpurun_out/r02_sanitizer_${TOOL}
mkdir -p gpurun_out
SEL=(
  "tests/test_gpu_decode.py::test_in_launch_merge_of_cut_rows"
  "tests/test_gpu_decode.py::test_fused_store_decode_equals_store_then_decode"
  "tests/test_gpu_decode.py::test_fused_store_with_foreign_slot_mapping"
  "tests/test_gpu_decode.py::test_overlap_mode_back_to_back"
  "tests/test_gpu_decode.py::test_decode_padding_rows_and_stale_tables"
  "tests/test_gpu_decode.py::test_decode_matches_oracle_and_flash_attn"
  "tests/test_gpu_cascade.py::test_cascade_fused_store_and_padding_rows"
  "tests/test_gpu_cascade.py::test_plan_kernel_matches_oracle"
  "tests/test_gpu_cascade.py::test_cascade_decode_matches_oracle_and_plain"
  "tests/test_gpu_prefill.py::test_prefill_matches_oracle_and_flash_attn"
  "tests/test_gpu_prefill.py::test_prefill_more_shapes_vs_oracle"
  "tests/test_gpu_prefill.py::test_warmup_path_without_cache"
  "tests/test_gpu_qkv_fused.py::test_normrope_store_vs_oracle"
  "tests/test_gpu_qkv_fused.py::test_qkv_decode_one_launch"
  "tests/test_gpu_store.py::test_store_edge_cases"
  "tests/test_gpu_store.py::test_store_matches_reference_golden"
  "tests/test_fast_meta.py::test_prepare_decode_matches_reference_builders"
  "tests/test_fast_meta.py::test_prepare_prefill_block256_batches"
)
export PYTORCH_NO_CUDA_MEMORY_CACHING=1
timeout ${SANITIZE_TIMEOUT:-1500} compute-sanitizer --tool "$TOOL" --error-exitcode 9 \
  --print-limit 20 --log-file "${OUT}.txt" \
  python -m pytest "${SEL[@]}" -q -p no:cacheprovider "$@" > "${OUT}_pytest.txt" 2>&1
RC=$?
echo "tool=$TOOL rc=$RC" | tee -a "${OUT}_pytest.txt"
tail -5 "${OUT}.txt"
tail -4 "${OUT}_pytest.txt"
exit 0
