# Final multi-GPU run of round 2:  gpurun --gpus N -- 'bash profiles/run_r02_final.sh N'
N=${1:-8}
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1 --nproc-per-node $N"
export IGG_DEVICE_TIMEOUT_S=20 IGG_HOST_TIMEOUT_S=60
t0=$(date +%s)
env IGG_MODE=auto IGG_DIRECT_THRESHOLD=0 IGG_BULK_THRESHOLD=16384 timeout 400 $TR --master-port 29541 tests/mp_worker.py > gpurun_out/mp${N}_auto_final_r02.log 2>&1
echo "mp$N auto rc=$? $(( $(date +%s) - t0 ))s: $(grep -c MISMATCH gpurun_out/mp${N}_auto_final_r02.log) mismatch lines; $(grep MP_WORKER_OK gpurun_out/mp${N}_auto_final_r02.log)"
timeout 300 $TR --master-port 29551 bench.py --gpus $N > gpurun_out/bench_r02_n$N.json 2> gpurun_out/bench_r02_n$N.err; echo "bench$N rc=$?"
python -c "
import json
d = json.loads(open('gpurun_out/bench_r02_n$N.json').read().strip().splitlines()[-1])
print($N, 'GPUs:', round(d['update_halo_us'], 2), 'us/call; per-call p50', round(d['per_call_us']['p50'], 2), 'graph', round(d['graph_replay_us_per_call'], 2), 'blocking', round(d['blocking_call_us']['p50'], 2), 'e2e', round(d['e2e']['us_per_step'], 1), 'nvl', d['nvlink_peer_copy_GBps'], d['parity']['golden_cases'], d['parity']['failed'], d['parity']['gather'], 'frac', round(d['roofline']['frac'], 3))" || tail -5 gpurun_out/bench_r02_n$N.err
timeout 300 $TR --master-port 29552 bench.py --gpus $N --impl reference > gpurun_out/bench_r02_n${N}_reference.json 2>/dev/null; cut -c1-200 gpurun_out/bench_r02_n${N}_reference.json
timeout 300 $TR --master-port 29564 profiles/bench_diffusion.py 256 200 > gpurun_out/diffusion_n${N}_final_r02.log 2>&1; echo "diffusion rc=$?"; grep '^{' gpurun_out/diffusion_n${N}_final_r02.log
if [ "$N" = "8" ]; then
  timeout 400 $TR --master-port 29561 profiles/bench_configs.py --modes ipc --steps 200 --only cfg2,cfg3,cfg4 --out gpurun_out/configs_r02_n8_final.jsonl > gpurun_out/configs_r02_n8_final.log 2>&1; echo "configs rc=$?"
  timeout 300 $TR --master-port 29562 profiles/bench_configs.py --modes ipc --steps 200 --sym --only "F=4" --out gpurun_out/configs_r02_n8_sym_final.jsonl > gpurun_out/configs_r02_n8_sym_final.log 2>&1; echo "configs sym rc=$?"
  for f in configs_r02_n8_final configs_r02_n8_sym_final; do python -c "
import json
for l in open('gpurun_out/$f.jsonl'):
    d = json.loads(l); print(d['config'][:64].ljust(64), d['mode'].ljust(6), d['update_halo_us'], 'us  roof', d['roofline_us'], 'frac', d['roofline_frac'])"; done
fi
