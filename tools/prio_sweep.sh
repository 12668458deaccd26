for p in "-3,-2,-2,0,-1" "-3,-2,-2,-1,0" "-3,-2,-2,-1,-1" "-3,-3,-3,0,-1" "-3,-2,-2,-2,0" "-3,-1,-1,-2,0" "-3,-2,-2,0,0"; do
  for rep in 1 2; do
    HF_CHAIN_PRIORITIES="$p" python bench.py --no-cpu-baseline --no-config5 --no-e2e 2>/dev/null | python -c "
import json,sys; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('$p', round(d['ms_per_step'],4), round(d['roofline']['ms_per_launch'],3))"
  done
done
