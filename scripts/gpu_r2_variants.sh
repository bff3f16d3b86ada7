# build compile-time variants of the sweep kernel on the GPU box (same image: nvcc is there), check parity, time at s = 1/2
# usage: bash scripts/gpu_r2_variants.sh "<defs 1>" "<defs 2>" ...   ("-" = defaults)
NZ=${NZ:-12}; NH=${NH:-48}
for defs in "$@"; do
  [ "$defs" = "-" ] && defs=""
  echo "#### defs: $defs"
  WXB_NVCC_DEFS="$defs" python wxfactory_b200/build.py --force > /dev/null || { echo BUILD FAILED; continue; }
  WXB_NO_REBUILD=1 timeout 600 python -m pytest tests/test_gpu_parity.py -m gpu -q -x -k "device_path or jvp_matches or host_pipeline" 2>&1 | tail -2
  WXB_NO_REBUILD=1 timeout 600 python bench.py --nz $NZ --nh $NH --steps 20 --warmup 5 --skip-cpu --skip-sw --skip-stepops --e2e-steps 3 2>&1 | tail -1 | python -c "
import json,sys
d=json.loads(sys.stdin.read())
for k in ('roofline','roofline_jvp'):
    r=d[k]; print(k, 'frac %.3f' % r['frac'], 'ms %.4f' % r['launch_ms']['evaluation'])
print('parity', d['parity']['ok'], d['parity']['rhs_rel_err'], d['parity']['jvp_err_over_dt_rhs_eps'], d['clocks']['sm_mhz'], d['clocks']['reasons'])
"
done
python wxfactory_b200/build.py --force > /dev/null
