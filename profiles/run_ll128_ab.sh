# A/B of the two IPC protocols on N GPUs:
#   gpurun --gpus N -- 'bash profiles/run_ll128_ab.sh N [full|quick|skip]'
# multi-rank parity worker in LL128 mode (full: all cases against the oracle; quick: golden vectors + full-size
# properties only), then bench.py with IGG_MODE=fence and IGG_MODE=ll128.
N=${1:-2}
PARITY=${2:-full}
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1"
export IGG_DEVICE_TIMEOUT_S=10 IGG_HOST_TIMEOUT_S=40
if [ "$PARITY" != "skip" ]; then
  [ "$PARITY" = "quick" ] && export IGG_WORKER_QUICK=1
  IGG_MODE=ll128 timeout 300 $TR --nproc-per-node $N --master-port 29541 tests/mp_worker.py > gpurun_out/mp${N}_ll128.log 2>&1; echo "mp$N ll128 ($PARITY) rc=$?"
  grep -v "^\*\|OMP_NUM\|^W1\|^$" gpurun_out/mp${N}_ll128.log | tail -8
  unset IGG_WORKER_QUICK
fi
for m in ${3:-fence ll128}; do
  IGG_MODE=$m timeout 150 $TR --nproc-per-node $N --master-port 2955$N bench.py --gpus $N --steps 200 --warmup 20 \
      > gpurun_out/bench_n${N}_$m.json 2> gpurun_out/bench_n${N}_$m.err; echo "bench$N $m rc=$?"
  python -c "
import json
d = json.loads(open('gpurun_out/bench_n${N}_$m.json').read().strip().splitlines()[-1])
print('$m', $N, 'GPUs:', round(d['ms_per_step'] * 1e3, 2), 'us', round(d['value'], 1), d['unit'], d['per_call_us'], d['parity'])" || tail -5 gpurun_out/bench_n${N}_$m.err
done
