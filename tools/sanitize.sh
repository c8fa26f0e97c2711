#!/bin/bash
# tools/sanitize.sh [outdir] -- compute-sanitizer over the hot-path kernels (memcheck / racecheck).
# Runs on a GPU box: memcheck and racecheck over three reference goldens through the C-ABI (c1 = the installation test,
# adv_ns_lower = rank/select path, adv_tandem = multi-distance first-wins), once more with the table restricted to mates,
# and memcheck over the multi-process peer-memory exchange.  One log per tool under outdir (default gpurun_out).
out=${1:-gpurun_out}
mkdir -p "$out"
cs=$(command -v compute-sanitizer || echo /usr/local/cuda/bin/compute-sanitizer)
sel='test_whole_path_matches_reference_goldens and (c1 or adv_ns_lower or adv_tandem)'
sel_m='test_mates_table_matches_reference_goldens and (adv_ns_lower or adv_tandem)'
export LINKS_EXCHANGE_TIMEOUT_MS=120000
for tool in memcheck racecheck; do
  timeout 1500 "$cs" --tool $tool --error-exitcode 9 --print-limit 20 python -m pytest tests/test_gpu_parity.py -x -q -k "$sel or $sel_m" \
    > "$out/sanitizer_$tool.log" 2>&1
  echo "exit code $?" >> "$out/sanitizer_$tool.log"
done
timeout 1500 "$cs" --tool memcheck --target-processes all --error-exitcode 9 --print-limit 20 python -m pytest tests/test_gpu_parity.py -x -q \
  -k "peer_memory_exchange" > "$out/sanitizer_memcheck_peer.log" 2>&1
echo "exit code $?" >> "$out/sanitizer_memcheck_peer.log"
grep -h -E "ERROR SUMMARY|exit code|passed|failed" "$out"/sanitizer_*.log
