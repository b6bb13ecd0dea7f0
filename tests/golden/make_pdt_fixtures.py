#!/usr/bin/env python
"""Draw the small .pdt populations committed under tests/golden/ (needs /root/reference's XMLs).

PLEDGE cannot run here (no JVM), so these populations come from featurenet_b200.sampler;
the files are in PLEDGE's format and are consumed by the reference's own Products_tree.py
in make_structural_golden.py and by the compat reader in the tests.
"""
import os
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))
from featurenet_b200 import sampler  # noqa: E402

REFERENCE = os.environ.get("FNET_REFERENCE", "/root/reference")

JOBS = [
    # (xml, kwargs for parse, n products, seed, p_select, output)
    ("nas_1_1_10.xml", {}, 8, 0, 0.7, "nas_1_1_10_seed0.pdt"),
    ("nas_5_5_10.xml", {}, 10, 0, 0.8, "nas_5_5_10_seed0.pdt"),
    ("main_1block_nas.xml", dict(nb_cells=5, nb_blocks=1), 6, 0, 0.8, "main_1block_nas_5cells_seed0.pdt"),
    ("datasets/main_light.xml", {}, 6, 0, 0.6, "main_light_seed0.pdt"),
]

for xml, kw, n, seed, p, out in JOBS:
    fm = sampler.parse_splot(os.path.join(REFERENCE, xml), **kw)
    products = sampler.sample_products(fm, n, seed=seed, p_select=p)
    assert all(sampler.check_product(fm, prod) for prod in products)
    sampler.write_pdt(os.path.join(HERE, out), fm, products)
    print(out, len(fm.names), "features", len(products), "products",
          [sum(1 for x in prod if x > 0) for prod in products])
