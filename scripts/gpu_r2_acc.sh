# accuracy A/B of the table-driven math in the sweep kernel: per-variable parity margins of every golden case
for defs in "$@"; do
  [ "$defs" = "-" ] && defs=""
  echo "#### defs: $defs"
  WXB_NVCC_DEFS="$defs" python wxfactory_b200/build.py --force > /dev/null || { echo BUILD FAILED; continue; }
  WXB_NO_REBUILD=1 python scripts/gpu_errors.py 2>&1 | grep -E "dcmip31 |case5|galewsky|dcmip21 |dcmip31_n4"
done
python wxfactory_b200/build.py --force > /dev/null
