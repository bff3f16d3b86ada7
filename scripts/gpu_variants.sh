# build and time compile-time variants of the n=4 main kernel on the GPU box (same image: nvcc is there)
for defs in "$@"; do
  echo "#### $defs"
  WXB_NVCC_DEFS="$defs" python wxfactory_b200/build.py --force > /dev/null || exit 1
  bash scripts/gpu_exp.sh 0 | grep -v "^=="
done
