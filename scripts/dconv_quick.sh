#!/bin/bash
# the resident-tile kernel alone on the shapes VERDICT names (compare with gpurun_out/dconv_ab.txt for FNET_DCONV=0)
for shape in "128 32 32 64 64 3 3 1" "128 32 32 32 32 3 3 1" "128 64 32 24 192 3 1 1" "128 14 14 12 120 5 5 1" "128 32 32 24 24 3 3 1" "128 16 16 192 192 3 3 1"; do
  for mode in tf32x3 tf32; do
    echo "== conv $shape $mode"
    FNET_DCONV=${FNET_DCONV:-2} timeout 60 python scripts/one_op.py conv $shape $mode 10 2>&1 | grep -v wgrad
  done
done
