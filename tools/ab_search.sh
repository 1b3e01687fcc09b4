# A/B of search-kernel builds on the bench workload: `bash tools/ab_search.sh "" _variantA _variantB`
# (variants built with `make -C sqlite_muninn_b200/csrc variant VARIANT=x EXTRA=-D…`)
for v in "$@"; do
  so=sqlite_muninn_b200/csrc/libmuninn_cuda$v.so
  MUNINN_CUDA_SO=$PWD/$so python bench.py --ef ${EF:-64} --no-cpu-baseline --steps 50 --warmup 5 ${BENCH_ARGS:-} 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); r=d['roofline']
print('VARIANT [$v]', 'kernel_ms', r['kernel_ms'], 'frac', r['frac'], 'gather', r.get('random_row_gather_peak'), 'value', round(d['value']), 'e2e', round(d['e2e']['value']), 'recall', d['config']['recall_at_10'])"
done
