# bench.py, both arms, defaults, at N GPUs:  gpurun --gpus N -- 'bash profiles/run_r02_bench_only.sh N'
N=${1:-8}
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1 --nproc-per-node $N"
timeout 300 $TR --master-port 29551 bench.py --gpus $N > gpurun_out/bench_r02_n$N.json 2> gpurun_out/bench_r02_n$N.err; echo "bench$N rc=$?"
timeout 300 $TR --master-port 29552 bench.py --gpus $N --impl reference > gpurun_out/bench_r02_n${N}_reference.json 2>/dev/null
python -c "
import json
d = json.loads(open('gpurun_out/bench_r02_n$N.json').read().strip().splitlines()[-1])
r = json.loads(open('gpurun_out/bench_r02_n${N}_reference.json').read().strip().splitlines()[-1])
print($N, 'GPUs:', round(d['update_halo_us'], 2), 'us/call; graph', round(d['graph_replay_us_per_call'], 2), 'p50', round(d['per_call_us']['p50'], 2), 'blocking', round(d['blocking_call_us']['p50'], 2), 'e2e', round(d['e2e']['us_per_step'], 1), 'cpu arm', round(r['update_halo_us'], 1), 'nvl', d['nvlink_peer_copy_GBps'], d['parity']['failed'], d['parity']['gather'], 'frac', round(d['roofline']['frac'], 3))" || tail -5 gpurun_out/bench_r02_n$N.err
