for mb in 32 64 128 256; do
  echo "== micro-batch $mb"
  timeout 300 python bench.py --steps 5 --warmup 3 --no-cpu --no-train --micro-batch $mb 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read()); print(round(d['value']), round(d['e2e']['value']), d['clocks']['sm_mhz'])"
done
