for mb in 128 256 128 64; do python bench.py --steps 5 --warmup 3 --no-cpu --no-train --micro-batch $mb 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print($mb, round(d['value']), round(d['e2e']['value']), d['clocks']['sm_mhz'])"; done
