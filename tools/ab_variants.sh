for v in variants_t512_c4 variants_t256_c5 variants_t1024_c5 variants_t256_c4 variants_t512_c6; do
  LINKS_GPU_LIB=$PWD/$v.so python bench.py --steps 5 --warmup 3 --no-cpu-baseline 2>/dev/null | python -c "
import json,sys; d=json.loads(sys.stdin.read()); print('$v', round(d['value']/1e9,3), round(d['ms_per_step'],3), 'extract', round(d['kernel_ms_per_step']['ms_extract'],3), 'e2e', round(d['e2e']['ms_per_step'],3))"
done
