# 1 GPU: big self-wraps (cfg4: 512^3 f32 hw 2 periodic, F fields) through the LL kernel (load/store), the fence kernel
# with the bulk-copy pipeline (IGG_BULK=1) and without (IGG_BULK=0); z faces only (16-byte aligned rows) and all dims.
for ud in 3 3,2 ""; do
  for cfg in "auto 1" "fence 1" "fence 0"; do
    set -- $cfg
    extra=""; [ -n "$ud" ] && extra="--updims $ud"
    IGG_MODE=$1 IGG_BULK=$2 timeout 200 python profiles/bench_configs.py --modes ipc --steps 100 --only cfg4 $extra 2>&1 | grep '^{' | python -c "
import sys, json
for l in sys.stdin:
    d = json.loads(l); print('IGG_MODE=$1 IGG_BULK=$2', d['config'][:56].ljust(56), d['update_halo_us'], 'us  roof', d['roofline_us'], 'frac', d['roofline_frac'])"
  done
done
