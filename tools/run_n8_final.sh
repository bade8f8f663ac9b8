# gpurun --gpus 8 -- bash tools/run_n8_final.sh : the multi-GPU lines of the final build (C3 views, C4 bands, C5 views; NCCL gather and peer-memory writes)
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29511"
Q="--no-batch-query --no-cpu-baseline --no-variants --no-gpu-reference"
$TR bench.py --gpus 8 $Q > gpurun_out/r02s_bench_C3_views_n8.json 2> gpurun_out/r02s_c3.err || tail -5 gpurun_out/r02s_c3.err
$TR bench.py --gpus 8 $Q --collective peer > gpurun_out/r02s_bench_C3_views_peer_n8.json 2> gpurun_out/r02s_c3p.err || tail -5 gpurun_out/r02s_c3p.err
$TR bench.py --gpus 8 $Q --workload C4 --sharding tiles > gpurun_out/r02s_bench_C4_tiles_n8.json 2> gpurun_out/r02s_c4.err || tail -5 gpurun_out/r02s_c4.err
$TR bench.py --gpus 8 $Q --workload C4 --sharding tiles --collective peer > gpurun_out/r02s_bench_C4_tiles_peer_n8.json 2> gpurun_out/r02s_c4p.err || tail -5 gpurun_out/r02s_c4p.err
$TR bench.py --gpus 8 $Q --workload C5 > gpurun_out/r02s_bench_C5_views_n8.json 2> gpurun_out/r02s_c5.err || tail -5 gpurun_out/r02s_c5.err
python -m pytest tests/test_gpu_peer_frames.py -q 2>&1 | tail -2
for f in C3_views C3_views_peer C4_tiles C4_tiles_peer C5_views; do python -c "
import json; d=json.loads(open('gpurun_out/r02s_bench_${f}_n8.json').read().strip().splitlines()[-1]); print('$f', 'ms/step', round(d['ms_per_step'],3), 'value %.4e'%d['value'], 'e2e ms/frame', round(d['e2e']['ms_per_frame'],3), [round(x,2) for x in d.get('per_rank_kernel_ms',[])])"; done
