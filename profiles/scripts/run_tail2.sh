for cfgv in "1 2" "2 2" "2 4" "3 3" "4 2" "4 4"; do set -- $cfgv; div=$1; wv=$2
  for cfgb in "cfg5 3072 60" "cfg5 384 300" "cfg3 48 300"; do set -- $cfgb
    DH_G_TAIL_DIV=$div DH_G_TAIL_WAVES_X2=$wv python bench.py --workload $1 --batch $2 --steps $3 --warmup 10 --no-cpu-baseline --no-e2e 2>/dev/null | python -c "
import json,sys; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('div $div waves_x2 $wv', '$1 B=$2', 'ms %.4f GB/s %.0f'%(d['ms_per_step'], d['roofline']['achieved']))"
  done; done
