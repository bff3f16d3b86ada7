# usage: gpu_variants_n.sh <order> "<defs>" ["<defs>" ...]   -- build + time variants of one order
n=$1; shift
for defs in "$@"; do
  WXB_NVCC_DEFS="$defs" python wxfactory_b200/build.py --force > /dev/null || { echo "build failed: $defs"; continue; }
  python scripts/sweep_orders.py $n quick 2>/dev/null | python -c "
import sys,json
for l in sys.stdin:
    d=json.loads(l); print('n=%d %-40s rhs %.3f ms (%.3f) jvp %.3f ms (%.3f) parity %.0e' % (d['n'], '$defs', d['rhs_ms'], d['rhs_frac_hbm'], d['jvp_ms'], d['jvp_frac_hbm'], d['parity_rhs']))
"
done
