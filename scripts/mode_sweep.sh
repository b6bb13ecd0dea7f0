#!/bin/bash
# fp32 (FFMA) against tf32x3 (tensor core, fp32-equivalent) per layer shape: the data behind the AUTO dispatch rule
run() { for m in fp32 tf32x3; do echo "== $* $m"; timeout 120 python scripts/one_op.py "$@" $m 5 2>&1 | tail -4 | head -3; done; }
run conv 128 14 14 12 120 5 5 1
run conv 128 28 28 6 12 5 5 1
run conv 128 32 32 3 32 3 3 1
run conv 128 32 32 32 32 3 3 1
run conv 128 16 16 32 64 3 3 1
run conv 128 16 16 64 64 3 3 1
run conv 128 32 32 24 24 3 3 1
run conv 128 16 16 48 96 3 1 1
run conv 128 8 8 96 192 5 5 2
run conv 128 32 32 16 16 1 1 1
run conv 128 32 32 8 64 1 1 1
run dense 25088 120 84
run dense 32768 84 120
run dense 8192 256 256
