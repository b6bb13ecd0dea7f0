#!/usr/bin/env python
"""Draw the benchmark populations shipped in this directory (needs the reference's feature models).

The reference's large populations (datasets/100Products.pdt, 1000Products.pdt of BASELINE.json's configs) are not in
its tree and PLEDGE cannot run without a JVM, so they are drawn here with featurenet_b200.sampler from the reference's
own nas_5_5_10.xml, in PLEDGE's .pdt format:

  nas_5_5_10_pop32_seed7.pdt         32 products, seed 7   (the default population of bench.py)
  nas_5_5_10_100Products_seed0.pdt   100 products, seed 0  (BASELINE configs[3]: `bench.py --population 100`)

  python featurenet_b200/data/make_populations.py [/root/reference]
"""
import os
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))
from featurenet_b200 import sampler  # noqa: E402

reference = sys.argv[1] if len(sys.argv) > 1 else "/root/reference"
fm = sampler.parse_splot(os.path.join(reference, "nas_5_5_10.xml"))
for count, seed, name in ((100, 0, "nas_5_5_10_100Products_seed0.pdt"),):
    products = sampler.sample_products(fm, count, seed=seed, p_select=0.8)
    assert len(products) == count and all(sampler.check_product(fm, p) for p in products)
    sampler.write_pdt(os.path.join(HERE, name), fm, products)
    print(name, len(fm.names), "features", len(products), "products")
