#!/bin/bash
# gathered (cp.async) against TMA-fed kernels on k x k stride-1 convolutions, tf32x3
run() { for t in 0 1; do echo "== $* FNET_TMA_CONV=$t"; FNET_TMA_CONV=$t timeout 120 python scripts/one_op.py "$@" tf32x3 5 2>&1 | tail -3; done; }
run conv 128 64 32 24 192 3 1 1
run conv 128 32 32 32 32 3 3 1
run conv 128 16 16 64 64 3 3 1
run conv 128 32 32 24 24 3 3 1
run conv 128 8 8 96 192 5 5 1
