# build and time compile-time variants of the shallow-water kernels on the GPU box
for defs in "$@"; do
  echo "#### $defs"
  WXB_NVCC_DEFS="$defs" python wxfactory_b200/build.py --force > /dev/null || exit 1
  timeout 300 python scripts/sw_bench.py galewsky ${SW_NH:-960} 2>&1 | tail -1 | python -c "
import json,sys
d=json.loads(sys.stdin.read())
print('rhs %.3f ms %s frac %.3f | jvp %.3f ms %s frac %.3f | parity %.1e %.1e' % (d['rhs_ms'], [round(x,3) for x in d['rhs_kernels_ms']], d['rhs_frac_hbm'], d['jvp_ms'], [round(x,3) for x in d['jvp_kernels_ms']], d['jvp_frac_hbm'], d['parity_rhs'], d['parity_jvp']))
"
done
