#!/bin/bash
# A/B of a FULL library build against the in-tree one on one box: the whole GPU test suite with the variant, then C3 / C4 / C5 of both.
#   make -C fv-srn_b200/csrc B=build_x EXTRA=-D... TARGET=../../ab/x.so -j8 ; gpurun -- 'bash tools/ab_full.sh x.so'
V=$1
cp fv-srn_b200/libfvsrn_b200.so /tmp/lib_keep.so
line() { python bench.py --workload $1 --steps 10 --warmup 3 --no-cpu-baseline --no-batch-query --no-gpu-reference --no-variants > /tmp/ab.json 2>/tmp/ab.err || tail -3 /tmp/ab.err
  python -c "
import json,sys;l=json.loads(open('/tmp/ab.json').read().strip().splitlines()[-1]);print(sys.argv[1], sys.argv[2], round(l['ms_per_frame'],3), round(l['roofline']['frac'],4), (l.get('parity') or {}).get('rgba_max_abs'))" $1 $2; }
cp ab/$V fv-srn_b200/libfvsrn_b200.so
python -m pytest tests -q -m gpu 2>&1 | tail -15
for w in C3 C4 C5; do line $w $V; done
cp /tmp/lib_keep.so fv-srn_b200/libfvsrn_b200.so
for w in C3 C4 C5; do line $w base; done
