"""A handful of memory-bound kernels, three launches each, for `ncu --set full` (see profiles/README.md)."""
import runpy
import sys

sys.argv = [sys.argv[0], "1"]          # one repetition after the three warm-up calls of scripts/membound.py
runpy.run_path(__file__.replace("ncu_membound.py", "membound.py"), run_name="__main__")
