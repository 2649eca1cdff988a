# Multi-GPU confirmation run (gpurun --gpus N -- 'bash profiles/run_multi_gpu.sh [N]'): the multi-rank parity worker,
# then bench.py at N, N/2, ... 2 GPUs.  Results land in gpurun_out/.
N=${1:-8}
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1"
timeout 300 $TR --nproc-per-node $N --master-port 29511 tests/mp_worker.py > gpurun_out/mp$N.log 2>&1; echo "mp$N rc=$?"
tail -3 gpurun_out/mp$N.log
n=$N
while [ $n -ge 2 ]; do
  timeout 120 $TR --nproc-per-node $n --master-port 2952$n bench.py --gpus $n --steps 200 --warmup 20 \
      > gpurun_out/bench_n$n.json 2> gpurun_out/bench_n$n.err; echo "bench$n rc=$?"
  python -c "
import json
d = json.loads(open('gpurun_out/bench_n$n.json').read().strip().splitlines()[-1])
print($n, 'GPUs:', d['ms_per_step'] * 1e3, 'us', d['value'], d['unit'])"
  n=$((n / 2))
done
