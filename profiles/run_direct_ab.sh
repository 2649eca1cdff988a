# Direct-to-halo vs staged, N GPUs in a row along z (z-faces cross NVLink):  bash profiles/run_direct_ab.sh N
N=${1:-2}
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1"
for sym in "" "--sym"; do
  for ud in "--updims 3" ""; do
    timeout 300 $TR --nproc-per-node $N --master-port 29581 profiles/bench_configs.py --modes ll128,fence --steps 200 --dims 1,1,$N $ud $sym --only cfg2,cfg4 2>&1 | grep '^{' | python -c "
import sys, json
for l in sys.stdin:
    d = json.loads(l); print(d['config'][:72].ljust(72), d['mode'].ljust(6), d['update_halo_us'], 'us  roof', d['roofline_us'], 'frac', d['roofline_frac'])"
  done
done
