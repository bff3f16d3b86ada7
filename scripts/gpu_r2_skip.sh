# where the sweep kernel's time goes: timing builds that leave one phase out (results are wrong; parity not checked)
NZ=${NZ:-12}; NH=${NH:-48}
for defs in "$@"; do
  [ "$defs" = "-" ] && defs=""
  echo "#### defs: $defs"
  WXB_NVCC_DEFS="$defs" python wxfactory_b200/build.py --force > /dev/null || { echo BUILD FAILED; continue; }
  WXB_NO_REBUILD=1 timeout 600 python bench.py --nz $NZ --nh $NH --steps 20 --warmup 5 --repeats 1 --sustain-s 0 --skip-check --skip-sweep --skip-cpu --skip-sw --skip-stepops --e2e-steps 3 2>&1 | tail -1 | python -c "
import json,sys
d=json.loads(sys.stdin.read())
for k in ('roofline','roofline_jvp'):
    r=d[k]; print(k, 'ms %.4f' % r['launch_ms']['evaluation'])
"
done
python wxfactory_b200/build.py --force > /dev/null
