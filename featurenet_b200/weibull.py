"""Reverse-Weibull location fits of CLEVER (model/metrics.py:316-317), optionally spread over worker processes.

The reference fits one distribution per (image, target class) with scipy's Nelder-Mead -- about 14 ms of pure
Python each, 4 500 of them for the 500-image robustness set, which is where the wall time of the CLEVER score goes
once the gradients come from the GPU (0.3 ms per image).  The fits are independent and deterministic, so they are
farmed out; every worker runs the very same scipy call, the results are identical to the serial loop.

This module imports numpy / scipy only: the workers never touch CUDA.
"""
import multiprocessing
import os

import numpy as np


def location(maxima, c_init=1):
    from scipy.optimize import fmin
    from scipy.stats import weibull_min
    return float(weibull_min.fit(-np.asarray(maxima, np.float64), c_init, optimizer=fmin)[1])


def _job(args):
    return location(*args)


def locations(all_maxima, c_init=1, workers=None):
    """One location per row of ``all_maxima``.  workers: None = FNET_CLEVER_WORKERS or min(cores, 16); <= 1 = serial."""
    if workers is None:
        workers = int(os.environ.get("FNET_CLEVER_WORKERS", min(os.cpu_count() or 1, 16)))
    jobs = [(list(map(float, m)), c_init) for m in all_maxima]
    if workers <= 1 or len(jobs) < 4 * workers:
        return [location(*j) for j in jobs]
    # "forkserver", never "fork": the caller is a multi-threaded process that owns a CUDA context, and a forked child
    # inherits locks held by threads that do not exist in it.  The server process is started once (from a clean
    # interpreter) and only ever imports this module: numpy and scipy, no CUDA.
    ctx = multiprocessing.get_context("forkserver")
    ctx.set_forkserver_preload(["numpy", "scipy.optimize", "scipy.stats"])
    with ctx.Pool(workers) as pool:
        return pool.map(_job, jobs, chunksize=max(1, len(jobs) // (8 * workers)))
