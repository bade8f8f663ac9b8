for rep in 1 2; do for pat in 4444 6532 543211 3333211 65311; do
FVSRN_HOST_BANDS=$pat python bench.py --steps 20 --warmup 3 --no-cpu-baseline --no-gpu-reference --no-variants --no-batch-query 2>/dev/null | python -c "
import json,sys; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('$pat', 'kernel %.3f'%d['ms_per_step'], {k:round(v['ms_per_step'],3) for k,v in d['e2e']['by_format'].items()})"
done; done
