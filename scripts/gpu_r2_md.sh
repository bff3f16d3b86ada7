#!/bin/bash
# single-pass multi-dot A/B (stepops.cu multidot_allrows): tests, then timing with and without it
export WXB_NO_REBUILD=1
timeout 600 python -m pytest tests/test_gpu_stepops.py tests/test_zz_no_alloc.py -m gpu -x -q 2>&1 | tail -5
for n in 53084160 424673280; do
  for op in 1 0; do
    echo "n=$n onepass=$op"
    WXB_VEC_ONEPASS=$op timeout 300 python scripts/vec_bench.py $n
  done
done
