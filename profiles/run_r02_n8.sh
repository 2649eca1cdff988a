# The 8-GPU run of round 2:  gpurun --gpus 8 -- 'bash profiles/run_r02_n8.sh [stress calls]'
N=8
STRESS=${1:-1000000}
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1 --nproc-per-node $N"
bash profiles/run_r02_multi.sh $N full no 0
# every BASELINE config in the default mode, and the NCCL comparison path
timeout 600 $TR --master-port 29561 profiles/bench_configs.py --modes ipc,nccl --steps 200 --out gpurun_out/configs_r02_n8.jsonl > gpurun_out/configs_r02_n8.log 2>&1; echo "configs rc=$?"
# the same with the fields in symmetric allocations (direct-to-halo from 4 MiB per call)
timeout 400 $TR --master-port 29562 profiles/bench_configs.py --modes ipc --steps 200 --sym --only cfg2,cfg3,cfg4 --out gpurun_out/configs_r02_n8_sym.jsonl > gpurun_out/configs_r02_n8_sym.log 2>&1; echo "configs sym rc=$?"
# forced protocols on the big configs
timeout 400 $TR --master-port 29563 profiles/bench_configs.py --modes ll128,fence --steps 200 --only cfg3ii,cfg4 --out gpurun_out/configs_r02_n8_proto.jsonl > gpurun_out/configs_r02_n8_proto.log 2>&1; echo "configs proto rc=$?"
for f in configs_r02_n8 configs_r02_n8_sym configs_r02_n8_proto; do
  python -c "
import sys, json
for l in open('gpurun_out/$f.jsonl'):
    d = json.loads(l)
    if 'update_halo_us' in d: print(d['config'][:64].ljust(64), d['mode'].ljust(6), d['update_halo_us'], 'us  roof', d['roofline_us'], 'frac', d['roofline_frac'])
    else: print(d['config'][:64].ljust(64), d.get('gather_ms'), 'ms  roof', d.get('roofline_ms'), 'frac', d.get('roofline_frac'))"
done
# diffusion3D weak scaling
timeout 300 $TR --master-port 29564 profiles/bench_diffusion.py 256 200 > gpurun_out/diffusion_n8_r02.log 2>&1; echo "diffusion rc=$?"; grep '^{' gpurun_out/diffusion_n8_r02.log
# LL128 under stress
IGG_MODE=ll128 timeout 900 $TR --master-port 29565 profiles/ll128_stress.py $STRESS > gpurun_out/ll128_stress_n8_r02.log 2>&1; echo "stress rc=$?"
grep "LL128_STRESS" gpurun_out/ll128_stress_n8_r02.log
