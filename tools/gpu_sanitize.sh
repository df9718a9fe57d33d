#!/bin/bash
# compute-sanitizer over the CI-size GPU tests:
# memcheck (out-of-bounds / misaligned accesses), racecheck (shared-memory hazards: the per-warp
# candidate buffers, k7_sort's warp-synchronous bitonic sort, k2_insert's claim protocol is global
# memory and is covered by memcheck + the parity of the rows) and initcheck on the small cases.
# Usage (on the GPU box): bash tools/gpu_sanitize.sh <tag>
TAG=${1:-r02s}
OUT=gpurun_out/$TAG; mkdir -p "$OUT"
SEL="test_golden_toy_fasta or test_golden_tiny_cases or test_golden_mid or test_anchor_stride or test_serial_verifier or test_strand_mirror or test_heavy_entries or test_capacity_overflow or test_variable_lengths or test_self_reverse"
FILES="tests/test_gpu_parity.py tests/test_filters.py tests/test_graph.py"
RSEL="test_golden_toy_fasta or test_golden_tiny_cases or edge_s or cfg2_x or test_heavy_entries or test_self_reverse or test_device_graph_against or test_device_filter_options"
for tool in memcheck racecheck; do
  K="$SEL or device"; [ $tool = racecheck ] && K="$RSEL"
  timeout 700 compute-sanitizer --tool $tool --target-processes all --print-limit 40 --error-exitcode 9 \
    --log-file "$OUT/sanitizer_$tool.log" \
    python -m pytest $FILES -m gpu -q -x -k "$K" -p no:cacheprovider > "$OUT/sanitizer_${tool}_pytest.log" 2>&1
  echo "$tool exit $?" | tee -a "$OUT/sanitizer_${tool}_pytest.log"
  tail -4 "$OUT/sanitizer_${tool}_pytest.log"; tail -8 "$OUT/sanitizer_$tool.log"
done
