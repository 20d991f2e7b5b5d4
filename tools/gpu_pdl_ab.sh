# A/B of programmatic dependent launch (CTN_PDL=0/1), eager launches (graphs never use it)
mkdir -p gpurun_out
for pdl in 0 1 0 1; do
  echo "== CTN_PDL=$pdl"
  CTN_PDL=$pdl timeout 300 python tools/bench_configs.py cfg1 2>/dev/null | python -c "
import json,sys
for l in sys.stdin:
    d=json.loads(l); print('  ', d['config'][:48], round(d.get('ms', d.get('ms_per_chunk', 0)),4))"
  CTN_PDL=$pdl timeout 300 python bench.py --steps 5 --warmup 3 --no-cpu --no-train --graph 0 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read()); print('   bench eager mb64', round(d['value']), round(d['e2e']['value']), d['clocks']['sm_mhz'])"
  CTN_PDL=$pdl timeout 300 python bench.py --steps 5 --warmup 3 --no-cpu --no-train --graph 0 --micro-batch 8 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read()); print('   bench eager mb8', round(d['value']), round(d['e2e']['value']), d['clocks']['sm_mhz'])"
done
