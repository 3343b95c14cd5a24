python -m pytest tests/test_sw.py -m gpu -q -x 2>&1 | tail -1
for sched in default 0.06,0.2,0.3,0.3,0.14 0.05,0.15,0.25,0.25,0.2,0.1; do
  if [ "$sched" != default ]; then export SW_CHUNK_SCHED=$sched; fi
  python tools/sw_bench.py --no-cpu --steps 8 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().split(chr(10))[-1])
print('$sched', 'SW value %.0f e2e %.0f (%.3f, %.2f ms) same %s'%(d['value'], d['e2e']['value'], d['e2e']['value']/d['value'], d['e2e']['ms_per_step'], d['e2e']['matches_resident_run']))"
done
