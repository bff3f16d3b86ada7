# timing experiments: WXB_KERNEL_OPTS bit mask (1 skip faces, 2 skip nodes [wrong results], 4 no TMA staging, 8 L2 prefetch)
for d in ${@:-0 4 12}; do
echo "== WXB_KERNEL_OPTS=$d"
WXB_KERNEL_OPTS=$d timeout 600 python bench.py --steps 20 --warmup 3 --skip-cpu --skip-sw --skip-stepops --e2e-steps 3 2>&1 | tail -1 | python -c "
import json,sys
d=json.loads(sys.stdin.read())
for k in ('roofline','roofline_jvp'):
    r=d[k]; print(k, 'frac %.3f' % r['frac'], r['launch_ms'])
print(d['parity']['ok'], d['parity']['rhs_rel_err'], d['clocks'])
"
done
