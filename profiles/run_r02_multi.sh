# Round-2 multi-GPU run:  gpurun --gpus N -- 'bash profiles/run_r02_multi.sh N [parity: full|skip] [configs: yes|no]'
# 1. the multi-rank oracle suite (tests/mp_worker.py) in the three IPC modes, logs -> gpurun_out/mpN_<mode>_r02.log
# 2. bench.py at N GPUs (golden + gather parity inside, peer-copy peak), default mode
# 3. profiles/bench_configs.py per protocol (LL128 vs fence) on the BASELINE configs
N=${1:-2}
PARITY=${2:-full}
CONFIGS=${3:-yes}
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1"
export IGG_DEVICE_TIMEOUT_S=20 IGG_HOST_TIMEOUT_S=60
if [ "$PARITY" = "full" ]; then
  for m in auto ll128 fence; do
    extra=""; [ "$m" = "auto" ] && extra="IGG_BULK_THRESHOLD=16384"
    t0=$(date +%s)
    env IGG_MODE=$m IGG_DIRECT_THRESHOLD=0 $extra timeout 500 $TR --nproc-per-node $N --master-port 2954$N tests/mp_worker.py > gpurun_out/mp${N}_${m}_r02.log 2>&1
    echo "mp$N $m rc=$? $(( $(date +%s) - t0 ))s: $(grep -c MISMATCH gpurun_out/mp${N}_${m}_r02.log) mismatch lines; $(grep MP_WORKER_OK gpurun_out/mp${N}_${m}_r02.log)"
    grep -v "^\*\|OMP_NUM\|^W1\|^$" gpurun_out/mp${N}_${m}_r02.log | tail -4
  done
fi
if [ "$PARITY" = "full" ]; then
  env IGG_MODE=auto IGG_LL128_SELFTEST_FAIL=1 IGG_WORKER_QUICK=1 timeout 300 $TR --nproc-per-node $N --master-port 2953$N tests/mp_worker.py > gpurun_out/mp${N}_fallback_r02.log 2>&1
  echo "mp$N forced self-test failure rc=$?: $(grep MP_WORKER_OK gpurun_out/mp${N}_fallback_r02.log) / $(grep -c 'LL128 is not safe' gpurun_out/mp${N}_fallback_r02.log) warning(s)"
fi
timeout 300 $TR --nproc-per-node $N --master-port 2955$N bench.py --gpus $N --steps 400 --warmup 50 \
    > gpurun_out/bench_n${N}_r02.json 2> gpurun_out/bench_n${N}_r02.err; echo "bench$N rc=$?"
python -c "
import json
d = json.loads(open('gpurun_out/bench_n${N}_r02.json').read().strip().splitlines()[-1])
print($N, 'GPUs:', round(d['ms_per_step'] * 1e3, 2), 'us/call; per-call p50', d['per_call_us']['p50'], 'graph', d['graph_replay_us_per_call'], 'blocking', d['blocking_call_us']['p50'], 'e2e', d['e2e']['us_per_step'], 'nvl', d['nvlink_peer_copy_GBps'], d['parity']['golden_cases'], d['parity']['failed'], d['parity']['gather'], d['roofline']['frac'])" || tail -5 gpurun_out/bench_n${N}_r02.err
if [ "$CONFIGS" = "yes" ]; then
  timeout 600 $TR --nproc-per-node $N --master-port 2956$N profiles/bench_configs.py --modes ll128,fence,nccl --steps 200 \
      --only cfg2,cfg3,cfg4 --out gpurun_out/configs_n${N}_r02.jsonl > gpurun_out/configs_n${N}_r02.log 2>&1; echo "configs rc=$?"
  grep '^{' gpurun_out/configs_n${N}_r02.log | python -c "
import sys, json
for l in sys.stdin:
    d = json.loads(l); print(d['config'][:48].ljust(48), d['mode'].ljust(6), d['update_halo_us'], 'us  roof', d['roofline_us'], 'frac', d['roofline_frac'])"
  if [ "$N" = "2" ]; then
    timeout 400 $TR --nproc-per-node $N --master-port 2957$N profiles/bench_configs.py --modes ll128,fence --steps 200 --dims 1,1,2 \
        --only cfg2,cfg3ii,cfg4 --out gpurun_out/configs_n${N}_z_r02.jsonl > gpurun_out/configs_n${N}_z_r02.log 2>&1; echo "configs z-split rc=$?"
    grep '^{' gpurun_out/configs_n${N}_z_r02.log | python -c "
import sys, json
for l in sys.stdin:
    d = json.loads(l); print('z-split', d['config'][:40].ljust(40), d['mode'].ljust(6), d['update_halo_us'], 'us  roof', d['roofline_us'], 'frac', d['roofline_frac'])"
  fi
fi
STRESS=${4:-0}
if [ "$STRESS" != "0" ]; then
  IGG_MODE=ll128 timeout 900 $TR --nproc-per-node $N --master-port 2958$N profiles/ll128_stress.py $STRESS > gpurun_out/ll128_stress_n${N}_r02.log 2>&1; echo "stress rc=$?"
  grep -v "^\*\|OMP_NUM\|^W1\|^$" gpurun_out/ll128_stress_n${N}_r02.log | tail -4
fi
