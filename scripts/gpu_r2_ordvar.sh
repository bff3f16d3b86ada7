# per-order variants of the sweep kernel: bash scripts/gpu_r2_ordvar.sh <n> "<defs>" ...   (10 M nodes, parity checked)
N=$1; shift
for defs in "$@"; do
  [ "$defs" = "-" ] && defs=""
  echo "#### n=$N defs: $defs"
  WXB_NVCC_DEFS="$defs" python wxfactory_b200/build.py --force > /dev/null 2>&1 || { echo BUILD FAILED; continue; }
  WXB_NO_REBUILD=1 timeout 600 python scripts/sweep_orders.py $N quick 2>/dev/null | tail -1 | python -c "
import json,sys
d=json.loads(sys.stdin.read())
print('rhs %.3f ms (%.3f)  jvp %.3f ms (%.3f)  parity %.1e %.1e' % (d['rhs_ms'], d['rhs_frac_hbm'], d['jvp_ms'], d['jvp_frac_hbm'], d['parity_rhs'], d['parity_jvp']))
"
done
python wxfactory_b200/build.py --force > /dev/null 2>&1
