# A/B of the host entry point on one box: staged bands (FVSRN_HOST_STAGED=1) vs direct writes into pinned host memory
for rep in 1 2; do for mode in staged direct; do
if [ $mode = staged ]; then export FVSRN_HOST_STAGED=1; else unset FVSRN_HOST_STAGED; fi
python bench.py --steps 20 --warmup 3 --no-cpu-baseline --no-gpu-reference --no-variants --no-batch-query 2>/dev/null | python -c "
import json,sys; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('$mode', 'kernel %.3f'%d['ms_per_step'], {k:round(v['ms_per_step'],3) for k,v in d['e2e']['by_format'].items()})"
done; done
