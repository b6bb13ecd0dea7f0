#!/bin/bash
# A/B of the resident-tile convolution kernel (csrc/dconv.cu) against the shifted-box TMA kernel it replaces, per shape
for shape in "128 32 32 64 64 3 3 1" "128 32 32 32 32 3 3 1" "128 32 32 32 64 3 3 1" "128 64 32 24 192 3 1 1" "128 14 14 12 120 5 5 1" "128 32 32 24 24 3 3 1" "128 16 16 24 24 3 3 1" "128 16 16 192 192 3 3 1"; do
  for mode in tf32x3 tf32; do
    for d in 0 2; do
      echo "== conv $shape $mode FNET_DCONV=$d"
      FNET_DCONV=$d timeout 60 python scripts/one_op.py conv $shape $mode 10 2>&1 | grep -v wgrad
    done
  done
done
