python -m pytest tests -m gpu -x -q 2>&1 | tail -3
for div in 1 2 4 8; do for wv in 2 3 5; do
  [ $div = 1 ] && [ $wv != 3 ] && continue
  for cfgb in "cfg3 48" "cfg5 384" "cfg2 16"; do set -- $cfgb
    DH_G_TAIL_DIV=$div DH_G_TAIL_WAVES_X2=$wv python bench.py --workload $1 --batch $2 --steps 300 --warmup 10 --no-cpu-baseline --no-e2e 2>/dev/null | python -c "
import json,sys; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('div $div waves_x2 $wv', '$1', 'ms %.4f GB/s %.0f'%(d['ms_per_step'], d['roofline']['achieved']))"
  done; done; done
