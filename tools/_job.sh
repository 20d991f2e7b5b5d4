python bench.py --steps 10 --warmup 3 --no-cpu 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print(round(d['value']), d['clocks']['sm_mhz'])
t=d['train_step']; print('train', t['ms_per_step'], t['gpu_launches_per_step'], t.get('serial_kernel_ms'))
for k,v in sorted(t['kernel_breakdown'].items(), key=lambda kv:-kv[1]['ms']): print('   ',k,v['launches'],round(v['ms']*1e3/v['launches'],1),'us',round(v['share'],3))"
