# time the bench's device step under different environment settings: bash scripts/gpu_env.sh "WXB_ORDER=0" "WXB_ORDER=1"
for e in "$@"; do
echo "== $e"
env $e timeout 600 python bench.py --steps 20 --warmup 3 --skip-cpu --skip-sw --skip-stepops --e2e-steps 3 2>&1 | tail -1 | python -c "
import json,sys
d=json.loads(sys.stdin.read())
for k in ('roofline','roofline_jvp'):
    r=d[k]; print(k, 'frac %.3f' % r['frac'], r['launch_ms'])
print(d['parity']['ok'], d['parity']['rhs_rel_err'], d['clocks'])
"
done
