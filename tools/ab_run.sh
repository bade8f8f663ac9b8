#!/bin/bash
# A/B of kernel builds on ONE B200 box (run-to-run variation between boxes is larger than most tuning steps):
#   tools/ab_build.sh <name> [-DFLAG..] builds each variant into ab/<name>.so (ab/ is git-ignored but travels with gpurun), then
#   gpurun -- 'tools/ab_run.sh C3 a.so b.so'
# Each variant is benchmarked twice, interleaved; prints workload, library, ms/frame, fraction of tensor peak, registers,
# and once per variant its RGBA max-abs against the reference's own kernel on a 512^2 frame (tools/diag_exact.py).
W=$1; shift
cp fv-srn_b200/libfvsrn_b200.so /tmp/lib_keep.so
for rep in 1 2; do
for l in "$@"; do
  cp ab/$l fv-srn_b200/libfvsrn_b200.so
  timeout 200 python bench.py --workload $W --steps 10 --warmup 3 --no-cpu-baseline --no-batch-query --no-gpu-reference > /tmp/ab.json 2>/tmp/ab.err || tail -3 /tmp/ab.err
  python -c "
import json,sys;l=json.load(open('/tmp/ab.json'));print(sys.argv[1], sys.argv[2], round(l['ms_per_frame'],3), round(l['roofline']['frac'],4), l['config']['kernel']['regs'])" $W $l
  if [ $rep = 1 ]; then case $W in C3*) timeout 200 python tools/diag_exact.py ${PARITY_CASE:-C3_relu_nerf} 512 tcgen05 2>&1 | grep -o "max-abs [0-9.e+-]*" ;; esac; fi
done; done
cp /tmp/lib_keep.so fv-srn_b200/libfvsrn_b200.so
