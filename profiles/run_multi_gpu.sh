set -x
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1"
timeout 240 $TR --nproc-per-node 8 --master-port 29511 tests/mp_worker.py > gpurun_out/mp8.log 2>&1; echo "mp8 rc=$?"
tail -3 gpurun_out/mp8.log
for n in 8 4 2; do
  timeout 120 $TR --nproc-per-node $n --master-port 2952$n bench.py --gpus $n --steps 200 --warmup 20 > gpurun_out/bench_n$n.json 2> gpurun_out/bench_n$n.err; echo "bench$n rc=$?"
  IGG_NO_ROTATE=1 timeout 120 $TR --nproc-per-node $n --master-port 2953$n bench.py --gpus $n --steps 200 --warmup 20 > gpurun_out/bench_n${n}_norot.json 2> gpurun_out/bench_n${n}_norot.err; echo "norot$n rc=$?"
  python - <<PY
import json
for f in ("gpurun_out/bench_n$n.json","gpurun_out/bench_n${n}_norot.json"):
    try:
        d=json.loads(open(f).read().strip().splitlines()[-1]); print(f, d["ms_per_step"], d["value"])
    except Exception as e: print(f, "ERR", e)
PY
done
