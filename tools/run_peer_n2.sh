# gpurun --gpus 2 -- bash tools/run_peer_n2.sh : PeerFrames test and the A/B NCCL gather vs peer-memory writes at N = 2
N=${N:-2}
python -m pytest tests/test_gpu_peer_frames.py -x -q > gpurun_out/r02l_peer_test.log 2>&1; tail -3 gpurun_out/r02l_peer_test.log
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511"
Q="--no-batch-query --no-cpu-baseline --no-variants --no-gpu-reference"
for c in gather peer; do
  $TR bench.py --gpus $N $Q --collective $c > gpurun_out/r02l_bench_C3_views_${c}_n$N.json 2> gpurun_out/r02l_c3_$c.err || tail -5 gpurun_out/r02l_c3_$c.err
  $TR bench.py --gpus $N $Q --workload C4 --sharding tiles --collective $c > gpurun_out/r02l_bench_C4_tiles_${c}_n$N.json 2> gpurun_out/r02l_c4_$c.err || tail -5 gpurun_out/r02l_c4_$c.err
  $TR bench.py --gpus $N $Q --workload C5 --collective $c > gpurun_out/r02l_bench_C5_views_${c}_n$N.json 2> gpurun_out/r02l_c5_$c.err || tail -5 gpurun_out/r02l_c5_$c.err
done
for f in C3_views_gather C3_views_peer C4_tiles_gather C4_tiles_peer C5_views_gather C5_views_peer; do python -c "
import json; d=json.loads(open('gpurun_out/r02l_bench_${f}_n$N.json').read().strip().splitlines()[-1]); print('$f', 'ms/step', round(d['ms_per_step'],3), 'e2e ms/frame', round(d['e2e']['ms_per_frame'],3), [round(x,2) for x in d.get('per_rank_kernel_ms',[])])"; done
