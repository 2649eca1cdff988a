TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1"
for b in 1 0; do
  IGG_BULK=$b timeout 300 $TR --nproc-per-node 2 --master-port 29571 profiles/bench_configs.py --modes fence,ll128 --steps 200 --dims 1,1,2 --updims 3 --only cfg4,cfg3ii > gpurun_out/zonly_bulk$b.log 2>&1
  echo "IGG_BULK=$b"; grep '^{' gpurun_out/zonly_bulk$b.log | python -c "
import sys, json
for l in sys.stdin:
    d = json.loads(l); print(d['config'][:60].ljust(60), d['mode'].ljust(6), d['update_halo_us'], 'us  roof', d['roofline_us'], 'frac', d['roofline_frac'])"
done
