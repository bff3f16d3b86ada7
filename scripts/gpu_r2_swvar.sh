# compile-time variants of the shallow-water tile kernel: bash scripts/gpu_r2_swvar.sh "<defs>" ...
for defs in "$@"; do
  [ "$defs" = "-" ] && defs=""
  echo "#### defs: $defs"
  WXB_NVCC_DEFS="$defs" python wxfactory_b200/build.py --force > /dev/null || { echo BUILD FAILED; continue; }
  WXB_NO_REBUILD=1 timeout 600 python scripts/sw_bench.py galewsky 960 2>&1 | tail -1 | python -c "
import json,sys
d=json.loads(sys.stdin.read())
print('rhs %.3f ms (%.3f)  jvp %.3f ms (%.3f)  parity %.1e %.1e' % (d['rhs_ms'], d['rhs_frac_hbm'], d['jvp_ms'], d['jvp_frac_hbm'], d['parity_rhs'], d['parity_jvp']))
"
done
python wxfactory_b200/build.py --force > /dev/null
