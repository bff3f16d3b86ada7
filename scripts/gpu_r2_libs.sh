# A/B of prebuilt library variants on the split path: bash scripts/gpu_r2_libs.sh lib1.so lib2.so ...
for lib in "$@"; do
for rep in 1 2; do
echo "== $lib ($rep)"
WXB_NO_REBUILD=1 WXB_LIB_PATH=$PWD/wxfactory_b200/$lib WXB_EULER_PATH=${EPATH:-split} timeout 600 python bench.py --nz 12 --nh 48 --steps 20 --warmup 5 --repeats 3 --sustain-s 0 --skip-sweep --skip-cpu --skip-sw --skip-stepops --e2e-steps 3 2>/dev/null | tail -1 | python -c "
import json,sys
d=json.loads(sys.stdin.read())
for k in ('roofline','roofline_jvp'):
    r=d[k]; print(k, 'frac %.3f' % r['frac'], r['launch_ms'])
print(d['parity']['ok'], d['clocks']['sm_mhz'], d['clocks']['reasons'])
"
done
done
