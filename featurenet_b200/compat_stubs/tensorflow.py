"""Name-only stand-in: /root/reference/full_evolution.py:13 and pledge_evolution.py:11 say ``import tensorflow`` and
never use the module.  featurenet_b200.install_compat(stubs=True) appends this directory to the END of sys.path, so
a real TensorFlow installation, if there is one, still wins."""
