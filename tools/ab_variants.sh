#!/bin/bash
# tools/ab_variants.sh build|run [workload] -- A/B of extract_kernel compile-time variants (LINKS_NVCC_DEFS) on one workload.
#   build : here (no GPU needed), variants/<name>.so, in parallel
#   run   : on the GPU box; one line per variant: job ms, extract ms, sectors issued per candidate
# variants/ is git-ignored and travels with gpurun.
set -u
cd "$(dirname "$0")/.."
declare -A V=(
  [base]=""
  [probe_with_set]="-DLINKS_PROBE_WITH_SET=1"
  [rest3]="-DLINKS_REST_FIRST=3"
  [rest3_pws]="-DLINKS_REST_FIRST=3 -DLINKS_PROBE_WITH_SET=1"
  [ctas5]="-DLINKS_EXTRACT_MIN_CTAS=5"
  [ctas3]="-DLINKS_EXTRACT_MIN_CTAS=3"
  [tile1024]="-DLINKS_EXTRACT_TILE=1024"
  [tile256]="-DLINKS_EXTRACT_TILE=256"
  [seq]="-DLINKS_SEQ_PROBES=1"
  [seq_ctas5]="-DLINKS_SEQ_PROBES=1 -DLINKS_EXTRACT_MIN_CTAS=5"
)
[ -n "${VARIANTS:-}" ] && names=$VARIANTS || names="${!V[@]}"
case "${1:-}" in
build)
  mkdir -p variants
  for n in $names; do
    ( nvcc -gencode arch=compute_100a,code=sm_100a -O3 -lineinfo -std=c++17 -Xcompiler -fPIC,-Wall ${V[$n]} -shared \
        -o variants/$n.so links_b200/csrc/links_gpu.cu 2> variants/$n.log && echo "built $n" || echo "FAILED $n" ) &
  done
  wait ;;
run)
  w=${2:-c3_celegans_100mbp}
  for n in $names; do
    LINKS_GPU_LIB=$PWD/variants/$n.so timeout 600 python bench.py --workload $w --steps ${STEPS:-3} --warmup 1 --no-cpu-baseline --no-e2e --no-extra 2>/dev/null | python -c "
import json,sys
try:
    d=json.loads(sys.stdin.read()); r=d['roofline']; s=r['issued_sectors_per_pair']
    print('$n', 'job_ms', round(d['ms_per_step'],2), 'extract_ms', round(d['kernel_ms_per_step']['ms_extract'],2), 'Gpairs/s', round(d['value']/1e9,2),
          'sectors/cand bloom', round(s['bloom'],2), 'set', round(s['member_set'],2), 'frac_rs', r.get('frac_random_sector'), 'digest', d['config']['rows_digest'])
except Exception as e:
    print('$n', 'FAILED', e)"
  done ;;
*) echo "usage: $0 build|run [workload]"; exit 2 ;;
esac
