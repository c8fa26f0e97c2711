#!/bin/bash
# tools/c4_ab.sh -- A/B of Bloom-build block / staging sizes and library variants on C4 at a tenth (one GPU).  MEASUREMENT SCRIPT.
# One line per arm: job ms, Bloom build ms, table build ms, extract ms, digest (must not move).
cd "$(dirname "$0")/.."
run() {  # name, env...
  local name=$1; shift
  env "$@" timeout 600 python bench.py --workload ${WORKLOAD:-c4_human_3gbp} --scale ${SCALE:-0.1} --steps ${STEPS:-2} --warmup 1 --no-cpu-baseline --no-e2e --no-extra 2>/dev/null | python -c "
import json,sys
try:
    d=json.loads(sys.stdin.read()); k=d['kernel_ms_per_step']
    print('$name', 'job_ms', round(d['ms_per_step'],1), 'bf', round(k['ms_bf_build'],1), 'table', round(k['ms_table_build'],1), 'extract', round(k['ms_extract'],1), 'classify', round(k['ms_classify'],1), 'digest', d['config']['rows_digest'], flush=True)
except Exception as e:
    print('$name', 'FAILED', e)"
}
for arm in "$@"; do
  case $arm in
    base) run base A=1 ;;
    block*) run $arm LINKS_BF_BLOCK_MB=${arm#block} ;;
    stage*) run $arm LINKS_BF_STAGE_MB=${arm#stage} ;;
    b32s*) run $arm LINKS_BF_BLOCK_MB=32 LINKS_BF_STAGE_MB=${arm#b32s} ;;
    lib:*) run $arm LINKS_GPU_LIB=$PWD/variants/${arm#lib:}.so ;;
    env:*) run $arm $(echo ${arm#env:} | tr "," " ") ;;
  esac
done
