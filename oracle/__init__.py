"""CPU oracle for the FeatureNet train+evaluate path.  TEST INFRASTRUCTURE ONLY; PARITY UNPINNED.

See oracle/ops.py for what that means and why.  Importers allowed: tests/, __graft_entry__.smoke(),
bench.py's cpu_baseline / --impl reference legs.
"""
