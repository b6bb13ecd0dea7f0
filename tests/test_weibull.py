"""The Weibull fits of CLEVER spread over worker processes give the numbers of the serial loop (CPU only)."""
import numpy as np

from featurenet_b200 import weibull
from oracle import attacks


def test_parallel_fits_equal_serial_and_the_oracle():
    rng = np.random.default_rng(4)
    maxima = [np.abs(rng.standard_normal(10)) + 1.0 + 0.1 * i for i in range(24)]
    serial = weibull.locations(maxima, 1, workers=1)
    spread = weibull.locations(maxima, 1, workers=3)
    assert serial == spread
    for m, loc in zip(maxima[:4], serial[:4]):
        assert loc == attacks.weibull_location(m, 1.0)
