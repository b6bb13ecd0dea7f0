"""featurenet_b200 -- B200-native train+evaluate engine for FeatureNet candidates.

The package holds exactly what the per-candidate hot path needs:

* ``csrc/``      hand-written sm_100a CUDA kernels behind the C ABI of ``include/fnet.h``
* ``lib``        ctypes binding of that C ABI (fails loudly when the .so is missing)
* ``layers``     symbolic layer constructors (the Keras surface the reference's
                 ``model/*.py`` calls) that record a layer DAG instead of a TF graph
* ``network``    executor: memory plan, CUDA-graph train/eval step, Keras ``fit`` semantics
* ``compat/``    modules with the reference's own names (``tensorflow_generator``,
                 ``helpers``, ``sgdr``, ``products_tree``, ``model.*``) so the reference
                 drivers run unchanged with ``PYTHONPATH=featurenet_b200/compat``

Nothing here imports ``oracle/``: that directory is test infrastructure.
"""
import os
import sys

__version__ = "0.1.0"

COMPAT_DIR = os.path.join(os.path.dirname(os.path.abspath(__file__)), "compat")


STUBS_DIR = os.path.join(os.path.dirname(os.path.abspath(__file__)), "compat_stubs")
BIN_DIR = os.path.join(COMPAT_DIR, "bin")


def install_compat(stubs=False, java_shim=False):
    """Put the reference-named modules (``tensorflow_generator``, ``model`` ...) on sys.path.

    stubs=True also makes ``import tensorflow`` resolve (two reference drivers import the name without using it);
    java_shim=True puts a ``java`` stand-in at the front of PATH that serves the one command the reference runs through
    the JVM, PLEDGE's ``generate_products`` (pledge_evolution.py:36-47), with featurenet_b200.sampler."""
    if COMPAT_DIR not in sys.path:
        sys.path.insert(0, COMPAT_DIR)
    if stubs and STUBS_DIR not in sys.path:
        sys.path.append(STUBS_DIR)
    if java_shim and BIN_DIR not in os.environ.get("PATH", "").split(os.pathsep):
        os.environ["PATH"] = BIN_DIR + os.pathsep + os.environ.get("PATH", "")
    return COMPAT_DIR
