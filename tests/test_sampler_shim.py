"""Population generation without PLEDGE: the sampler honours tree, group and cross-tree clauses,
including the negated-leaf constraints that PledgeEvolution.generate_children appends
(/root/reference/pledge_evolution.py:70-99), and the `java` stand-in serves the reference's own command line
(pledge_evolution.py:36-47)."""
import os
import shutil
import subprocess
import sys
import xml.etree.ElementTree as ET

import pytest

import featurenet_b200
from featurenet_b200 import sampler

HERE = os.path.dirname(os.path.abspath(__file__))
TINY = os.path.join(HERE, "data", "tiny_fm.xml")
REFERENCE = os.environ.get("FNET_REFERENCE_DIR", "/root/reference")


def selected(fm, signed):
    return {fm.names[i - 1] for i in signed if i > 0}


def test_every_sampled_product_satisfies_every_clause():
    fm = sampler.parse_splot(TINY)
    products = sampler.sample_products(fm, 40, seed=3)
    assert len(products) >= 8 and len({tuple(p) for p in products}) == len(products)
    for signed in products:
        assert sampler.check_product(fm, signed)
        names = selected(fm, signed)
        assert {"Root", "Architecture", "Stem"} <= names
        assert len(names & {"Stem_conv", "Stem_pool", "Stem_dense"}) == 1            # [1,1] group
        if "Extra" in names:
            assert names & {"Extra_bn", "Extra_drop"}                                # [1,*] group
        else:
            assert not names & {"Extra_bn", "Extra_drop"}
        assert "Stem_pool" not in names or "Extra" in names                          # C1
        assert not {"Skip", "Extra_drop"} <= names                                   # C2
    # check_product rejects what breaks a clause
    bad = [i if fm.names[i - 1] in ("Root", "Architecture", "Stem", "Stem_conv", "Stem_pool") else -i
           for i in range(1, len(fm.names) + 1)]
    assert not sampler.check_product(fm, bad)


def add_negated_leaves(src, dst, labels):
    """The edit generate_children makes to the XML (pledge_evolution.py:79-91), restated: one clause
    `~Architecture  or  ~<label>` per chosen leaf, appended to the constraints text."""
    tree = ET.parse(src)
    constraints = list(tree.getroot())[1]
    raw = constraints.text.split("\n")
    count = len(raw)
    raw = raw[:-1]
    for i, label in enumerate(labels):
        raw.append("C{}:~Architecture  or  ~{}".format(count + i - 1, label))
    constraints.text = "\n".join(raw)
    tree.write(dst, encoding="UTF-8", xml_declaration=True)


def test_negated_leaf_constraints_are_honoured(tmp_path):
    dst = str(tmp_path / "mutant.xml")
    add_negated_leaves(TINY, dst, ["Stem_conv", "Extra_bn"])
    fm = sampler.parse_splot(dst)
    products = sampler.sample_products(fm, 30, seed=5)
    assert products
    for signed in products:
        names = selected(fm, signed)
        assert sampler.check_product(fm, signed)
        assert "Stem_conv" not in names and "Extra_bn" not in names


def test_java_stand_in_serves_pledges_command_line(tmp_path):
    out = str(tmp_path / "products.pdt")
    env = dict(os.environ, PATH=featurenet_b200.BIN_DIR + os.pathsep + os.environ.get("PATH", ""))
    assert shutil.which("java", path=env["PATH"]) == os.path.join(featurenet_b200.BIN_DIR, "java")
    subprocess.check_call(["java", "-jar", "./PLEDGE.jar", "generate_products", "-fm", TINY, "-nbProds", "6", "-o", out,
                           "-timeAllowedMS", "1000"], env=env)
    from products_tree import ProductSet
    import contextlib
    import io
    with contextlib.redirect_stdout(io.StringIO()):
        loaded = ProductSet(out, binary_products=True)
    assert loaded.nbProducts == 6 and loaded.nbFeatures == 10
    rc = subprocess.call(["java", "-version"], env=env, stderr=subprocess.DEVNULL)
    assert rc == 64                                                                     # anything else is refused


@pytest.mark.skipif(not os.path.isfile(os.path.join(REFERENCE, "pledge_evolution.py")),
                    reason="the reference tree is only present in the build container")
def test_reference_generate_children_runs_without_a_jvm(tmp_path, monkeypatch):
    """The reference's OWN PledgeEvolution.generate_children / run_pledge, imported unmodified, with the stand-in on PATH."""
    featurenet_b200.install_compat(stubs=True, java_shim=True)
    monkeypatch.setattr(sys, "path", sys.path + [REFERENCE])      # AFTER compat/: only the driver itself comes from there
    monkeypatch.chdir(tmp_path)
    import importlib
    if not hasattr(ET.Element, "getchildren"):
        pytest.skip("pledge_evolution.py:76 calls Element.getchildren(), removed in Python 3.9: the reference itself "
                    "cannot run generate_children on this interpreter")
    pledge_evolution = importlib.import_module("pledge_evolution")
    out = pledge_evolution.PledgeEvolution.generate_children(str(tmp_path / "gen1"), ["Stem_conv"], TINY, 5, constraints_ratio=1.1)
    assert os.path.isfile(out)


@pytest.mark.skipif(not os.path.isfile(os.path.join(REFERENCE, "pledge_evolution.py")),
                    reason="the reference tree is only present in the build container")
def test_reference_run_pledge_runs_without_a_jvm(tmp_path, monkeypatch):
    featurenet_b200.install_compat(stubs=True, java_shim=True)
    monkeypatch.setattr(sys, "path", sys.path + [REFERENCE])
    import importlib
    pledge_evolution = importlib.import_module("pledge_evolution")
    out = str(tmp_path / "base.pdt")
    assert pledge_evolution.run_pledge(TINY, 4, out, pledge_evolution.PledgeEvolution.pledge_path, duration=1) == 0
    fm = sampler.parse_splot(TINY)
    with open(out) as handle:
        lines = [ln.strip() for ln in handle if ln.strip()]
    products = [[int(v) for v in ln.split(";") if v] for ln in lines if "->" not in ln]
    assert len(products) == 4 and all(sampler.check_product(fm, p) for p in products)
