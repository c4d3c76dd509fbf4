#!/bin/bash
# A/B of compile-time kernel variants on one box: build/variants/libnsof_<tag>.so (nvcc -D... of the same sources),
# each run through the bench command; prints value, full-frame value and the per-class kernel times.
# usage: scripts/variant_bench.sh A B C ...
for v in "$@"; do
  NSOF_LIB_PATH=$PWD/build/variants/libnsof_$v.so python bench.py --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/variant_$v.json 2> gpurun_out/variant_$v.err || { echo "$v failed"; tail -3 gpurun_out/variant_$v.err; continue; }
  python - "$v" <<'PY'
import json,sys
v=sys.argv[1]
d=json.loads(open(f'gpurun_out/variant_{v}.json').read().strip().splitlines()[-1])
ff=d['full_frame']; k=ff['kernels']
cls={}
for name,r in k.items():
    c=name.split('.')[0]; cls[c]=cls.get(c,0)+r['ms_per_step']
print(v, 'value %.0f e2e %.0f full_frame %.0f (%.3f of peak) roofline %.3f'%(d['value'],d['e2e']['value'],ff['value'],ff['frac_of_peak'],d['roofline']['frac']), ' '.join(f'{c}={t:.3f}' for c,t in cls.items()))
PY
done
