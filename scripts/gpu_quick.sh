# quick GPU check: parity tests, then a short bench line (no CPU arm)
timeout 900 python -m pytest tests -m gpu -q -x 2>&1 | tail -8
timeout 600 python bench.py --steps 20 --warmup 3 --skip-cpu --skip-sw --skip-stepops --e2e-steps 3 2>&1 | tail -1 | python -c "
import json,sys
d=json.loads(sys.stdin.read())
print('value %.3e  ms/step %.3f  e2e %.3e' % (d['value'], d['ms_per_step'], d['e2e']['value']))
for k in ('roofline','roofline_jvp'):
    r=d[k]; print(k, 'frac %.3f' % r['frac'], r['launch_ms'])
print(d['parity']['ok'], d['parity']['rhs_rel_err'], d['clocks'])
"
