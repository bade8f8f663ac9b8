TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29511"
$TR bench.py --gpus 8 --workload C4 --sharding tiles --steps 20 --warmup 3 > gpurun_out/r02i_bench_C4_tiles_n8.json 2> gpurun_out/r02i_c4.err
$TR bench.py --gpus 8 --workload C4 --sharding tiles --no-rebalance --steps 20 --warmup 3 > gpurun_out/r02i_bench_C4_tiles_norebalance_n8.json 2>> gpurun_out/r02i_c4.err
$TR bench.py --gpus 8 --workload C5 > gpurun_out/r02i_bench_C5_views_n8.json 2> gpurun_out/r02i_c5.err
$TR bench.py --gpus 8 > gpurun_out/r02i_bench_C3_views_n8.json 2> gpurun_out/r02i_c3.err
for f in C4_tiles_n8 C4_tiles_norebalance_n8 C5_views_n8 C3_views_n8; do python -c "
import json,sys; d=json.loads(open('gpurun_out/r02i_bench_${f}.json').read().strip().splitlines()[-1]); print('$f', round(d['ms_per_step'],3), '%.3e'%d['value'], [round(x,2) for x in d.get('per_rank_kernel_ms',[])], d['config']['sharding'][:160])" || tail -5 gpurun_out/r02i_c4.err; done
