"""Reader of PLEDGE product files (.pdt) -> nested block trees.

Behaviour of /root/reference/Products_tree.py:11-138 (the drivers import it as
``products_tree``; the reference ships the file as ``Products_tree.py``, which only resolves
on case-insensitive file systems -- this module is the lowercase name they ask for).

A .pdt file is a list of ``id->Label`` lines followed by one line per product holding
``;``-terminated signed feature ids.  ``format_product`` keeps the selected ids, hangs every
feature under the feature whose label is its own label minus the last ``_segment`` and
returns the top-level ``Block<k>`` trees.
"""
import os
import re


class ProductSetError(Exception):
    pass


_FEATURE_LINE = re.compile(r"(\d*)->(\w*)")


def _is_selected(token):
    return str(token).isdigit() and int(token) >= 0


class ProductSet(object):

    @staticmethod
    def filter_leaves(features, keep_list=None):
        # the reference computes a keep-list filter and then overwrites it (Products_tree.py:15-23)
        return {k: features[k] for k in features.keys() if len(k) > 3}

    def __init__(self, url, binary_products=False):
        self.features = {}
        self.constraints = []
        self.products = []
        self.nbFeatures = 0
        self.nbProducts = 0
        self.last_products_url = ""
        self.binary_products = binary_products
        if url:
            self.load_products_from_url(url)

    def load_products_from_url(self, url):
        print("==> Loading products from {}".format(url))
        if not os.path.isfile(url):
            print("File not found")
            raise ProductSetError()

        self.last_products_url = url
        self.features, self._features_reverse = {}, {}
        self.constraints, self.products = [], []
        with open(url, "r") as handle:
            for line in handle:
                declared = _FEATURE_LINE.match(line)
                if declared:
                    self.features[declared.group(1)] = declared.group(2)
                    self._features_reverse[declared.group(2)] = declared.group(1)
                    continue
                tokens = line.split(";")[:-1]
                if self.binary_products:
                    tokens = sorted(tokens, key=lambda k: abs(int(k)))
                    tokens = [1 if str(x).isdigit() and int(x) > 0 else 0 for x in tokens]
                self.products.append(tokens)
        self.nbFeatures = len(self.features)
        self.nbProducts = len(self.products)

    def format_product(self, prd_index=0, original_product=None, include_original=True, sort_features=False):
        original_product = self.products[prd_index] if not original_product else original_product
        if not original_product:
            return None

        if self.binary_products:
            selected = [i + 1 for i in range(len(original_product)) if int(original_product[i]) > 0]
        else:
            selected = [abs(int(x)) for x in original_product if _is_selected(x)]

        position = {self.features[str(fid)]: i for i, fid in enumerate(selected)}
        nodes = [{"label": self.features[str(fid)], "id": fid, "children": []} for fid in selected]

        # children are attached walking the ids backwards, so they end up in descending id order
        for fid in selected[::-1]:
            label = self.features.get(str(fid))
            cut = label.rfind("_")
            parent_label = label[0:cut] if cut > -1 else ""
            if parent_label:
                parent_at = position.get(parent_label)
                if parent_at:       # (sic) a parent at position 0 is skipped, as in the reference
                    nodes[parent_at].get("children").append(nodes[selected.index(fid)])

        blocks = [n for n in nodes if n.get("label").startswith("Block") and n.get("label").find("_") == -1]
        if not include_original:
            return blocks
        listed = sorted(original_product, key=lambda k: abs(int(k))) if sort_features else original_product
        return blocks, listed

    def light_product(self, prd_index=0, product=None):
        if not product:
            product = self.products[prd_index]
        trees, _ = self.format_product(original_product=product)

        def shorten(node):
            node["label"] = node["label"][node["label"].rfind("_"):]
            node["children"] = sorted([shorten(c) for c in node["children"]], key=lambda k: k["id"])
            return node

        return [shorten(t) for t in trees]

    def format_products(self, include_original=True, sort_features=False):
        return [self.format_product(original_product=p, include_original=include_original,
                                    sort_features=sort_features) for p in self.products]

    def light_products(self):
        for product in self.products:
            yield self.light_product(product=product)
