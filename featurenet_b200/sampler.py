"""Constraint-aware product sampler for FeatureNet feature models (.xml -> .pdt).

The reference samples products with PLEDGE (a Java/SAT4J jar run through ``java -jar``,
/root/reference/pledge_evolution.py:36-47).  There is no JVM on the build or GPU boxes and
the large ``.pdt`` populations named by BASELINE.json are not shipped, so benchmark
populations are drawn here instead.  Output is the PLEDGE file format read by
``products_tree.ProductSet`` (/root/reference/Products_tree.py:51-70): ``id->Label`` lines in
level order, then one line of ``;``-terminated signed ids per product.

This is a host-side convenience, not an accelerated path: a
randomised DPLL over the clauses of the SPLOT tree plus the cross-tree constraints.  It
makes no attempt at PLEDGE's similarity-driven diversity.
"""
import random
import re
import xml.etree.ElementTree as ET

_NODE = re.compile(r"^(\t*):([romg]?)\s*(.*)$")


class FeatureModel(object):
    def __init__(self):
        self.names = []          # index -> label, level (BFS) order, ids are index + 1
        self.index = {}
        self.clauses = []        # lists of signed 1-based ids
        self.optional = set()

    def var(self, name):
        return self.index[name] + 1


def expand_template(text_tree, text_constraints, nb_cells, nb_blocks):
    """Same expansion as /root/reference/extender.py:3-62 on the raw strings ([k] blocks, [i] cells)."""
    cell_at = text_tree.find(":o Block[k]_Element[i]")
    cell = text_tree[cell_at:]
    text_tree = text_tree[:cell_at] + "\t\t\t\t\t".join(cell.replace("[i]", str(i + 1)) for i in range(nb_cells))
    block_at = text_tree.find(":o Block[k](Block[k])")
    block = text_tree[block_at:]
    text_tree = text_tree[:block_at] + "\t\t\t\t".join(block.replace("[k]", str(k + 1)) for k in range(nb_blocks))

    out = []
    for raw in text_constraints.split("\n"):
        parts = raw.split(":")
        if len(parts) < 2:
            continue
        body = parts[1]
        if "[k]" not in body:
            out.append(body)
            continue
        if "CLC" in parts[0]:
            continue
        for k in range(nb_blocks):
            if "[k+1]" in body and k + 1 == nb_blocks:
                continue
            with_k = body.replace("[k]", str(k + 1)).replace("[k+1]", str(k + 2))
            if "[i]" not in body:
                out.append(with_k)
                continue
            for i in range(nb_cells):
                if "[i+1]" in body and i + 1 == nb_cells:
                    continue
                out.append(with_k.replace("[i]", str(i + 1)).replace("[i+1]", str(i + 2)))
    constraints = "\n".join("C%d:%s" % (n + 1, c) for n, c in enumerate(out))
    return text_tree, constraints


def parse_splot(path, nb_cells=None, nb_blocks=None):
    """Read a SPLOT/FeatureIDE-text feature model; expand [k]/[i] templates when sizes are given."""
    root = ET.parse(path).getroot()
    tree_text = root.find("feature_tree").text
    cons_text = root.find("constraints").text or ""
    if "[k]" in tree_text:
        tree_text, cons_text = expand_template(tree_text, cons_text, nb_cells or 1, nb_blocks or 1)

    # --- tree: indentation gives depth; ':g' nodes are anonymous XOR/OR groups
    entries = []                 # (depth, kind, name or cardinality)
    for line in tree_text.split("\n"):
        if not line.strip():
            continue
        m = _NODE.match(line)
        if not m:
            continue
        depth, kind, rest = len(m.group(1)), m.group(2), m.group(3).strip()
        if kind == "g":
            entries.append((depth, "g", rest))
        else:
            entries.append((depth, kind or "c", rest[:rest.find("(")] if "(" in rest else rest))

    class _N(object):
        def __init__(self, kind, name):
            self.kind, self.name, self.children, self.parent = kind, name, [], None

    top, stack = None, []
    for depth, kind, name in entries:
        node = _N(kind, name)
        while stack and stack[-1][0] >= depth:
            stack.pop()
        if stack:
            node.parent = stack[-1][1]
            stack[-1][1].children.append(node)
        else:
            top = node
        stack.append((depth, node))

    fm = FeatureModel()
    # level order over named features; ':g' group nodes are transparent
    fm.names, fm.index = [], {}
    level = [top]
    while level:
        nxt = []
        for node in level:
            fm.index[node.name] = len(fm.names)
            fm.names.append(node.name)
            for child in node.children:
                if child.kind == "g":
                    nxt.extend(child.children)
                else:
                    nxt.append(child)
        level = nxt

    def named_parent(node):
        p = node.parent
        while p is not None and p.kind == "g":
            p = p.parent
        return p

    def walk(node):
        if node.kind == "r":
            fm.clauses.append([fm.var(node.name)])
        for child in node.children:
            if child.kind == "g":
                owner = fm.var(named_parent(child).name)
                members = [fm.var(c.name) for c in child.children]
                lo, hi = re.findall(r"\[(\d+|\*),(\d+|\*)\]", child.name)[0]
                fm.clauses.append([-owner] + members)                  # at least one
                for mvar in members:
                    fm.clauses.append([-mvar, owner])
                if hi == "1":
                    for a in range(len(members)):
                        for b in range(a + 1, len(members)):
                            fm.clauses.append([-members[a], -members[b]])
                for c in child.children:
                    fm.optional.add(fm.var(c.name))
                    walk(c)
            else:
                me, parent = fm.var(child.name), fm.var(named_parent(child).name)
                fm.clauses.append([-me, parent])
                if child.kind == "m":
                    fm.clauses.append([-parent, me])
                else:
                    fm.optional.add(me)
                walk(child)

    walk(top)

    for raw in cons_text.split("\n"):
        if ":" not in raw:
            continue
        body = raw.split(":", 1)[1]
        clause = []
        for lit in body.split(" or "):
            lit = lit.strip()
            if not lit:
                continue
            neg = lit.startswith("~")
            name = lit[1:].strip() if neg else lit
            if name not in fm.index:
                clause = None
                break
            clause.append(-fm.var(name) if neg else fm.var(name))
        if clause:
            fm.clauses.append(clause)
    return fm


class _Solver(object):
    """Tiny DPLL with unit propagation; decisions follow level order with random polarity."""

    def __init__(self, fm, rng, p_select):
        self.fm, self.rng, self.p = fm, rng, p_select
        self.n = len(fm.names)
        self.occ = {}
        for ci, clause in enumerate(fm.clauses):
            for lit in clause:
                self.occ.setdefault(lit, []).append(ci)

    def _assign(self, value, trail, lit):
        """Set lit true; propagate. Returns False on conflict."""
        pending = [lit]
        while pending:
            cur = pending.pop()
            v = abs(cur)
            have = value[v]
            if have is not None:
                if have != (cur > 0):
                    return False
                continue
            value[v] = cur > 0
            trail.append(v)
            for ci in self.occ.get(-cur, ()):
                unassigned, satisfied = None, False
                count = 0
                for other in self.fm.clauses[ci]:
                    ov = value[abs(other)]
                    if ov is None:
                        unassigned = other
                        count += 1
                    elif ov == (other > 0):
                        satisfied = True
                        break
                if satisfied:
                    continue
                if count == 0:
                    return False
                if count == 1:
                    pending.append(unassigned)
        return True

    def solve(self, forced=()):
        value = [None] * (self.n + 1)
        trail = []
        for clause in self.fm.clauses:
            if len(clause) == 1 and not self._assign(value, trail, clause[0]):
                return None
        for lit in forced:
            if not self._assign(value, trail, lit):
                return None
        if not self._search(value, 1):
            return None
        return value

    def _search(self, value, start):
        v = start
        while v <= self.n and value[v] is not None:
            v += 1
        if v > self.n:
            return True
        first = self.rng.random() < self.p
        for pick in (first, not first):
            trail = []
            if self._assign(value, trail, v if pick else -v) and self._search(value, v + 1):
                return True
            for t in trail:
                value[t] = None
        return False


def sample_products(fm, count, seed=0, p_select=0.6, forced=()):
    """Draw ``count`` distinct valid products; each is a list of signed ids for every feature."""
    import sys
    rng = random.Random(seed)
    solver = _Solver(fm, rng, p_select)
    limit = sys.getrecursionlimit()
    sys.setrecursionlimit(max(limit, 4 * len(fm.names) + 1000))
    products, seen, attempts = [], set(), 0
    try:
        while len(products) < count and attempts < 50 * count + 100:
            attempts += 1
            value = solver.solve(forced)
            if value is None:
                break
            signed = tuple((i if value[i] else -i) for i in range(1, len(fm.names) + 1))
            if signed in seen:
                continue
            seen.add(signed)
            products.append(list(signed))
    finally:
        sys.setrecursionlimit(limit)
    return products


def check_product(fm, signed):
    chosen = set(x for x in signed if x > 0)
    for clause in fm.clauses:
        if not any((lit in chosen) if lit > 0 else (-lit not in chosen) for lit in clause):
            return False
    return True


def write_pdt(path, fm, products):
    with open(path, "w") as handle:
        for i, name in enumerate(fm.names):
            handle.write("%d->%s\n" % (i + 1, name))
        for signed in products:
            handle.write("".join("%d;" % x for x in signed) + "\n")
    return path


def generate_products(feature_model, nb_products, output, seed=None, nb_cells=None, nb_blocks=None, p_select=0.6):
    """What ``java -jar PLEDGE.jar generate_products -fm <xml> -nbProds <n> -o <pdt>`` does for the reference
    (pledge_evolution.py:36-47): read the feature model -- including the negated-leaf constraints that
    ``PledgeEvolution.generate_children`` appends (pledge_evolution.py:70-99, ``~Architecture  or  ~<leaf>``) --
    draw n valid products and write them in PLEDGE's .pdt format.  Returns the number written."""
    fm = parse_splot(feature_model, nb_cells=nb_cells, nb_blocks=nb_blocks)
    if seed is None:
        seed = random.SystemRandom().randrange(1 << 30)
    products = sample_products(fm, int(nb_products), seed=seed, p_select=p_select)
    if not products:
        raise SystemExit("the feature model %s has no valid product" % feature_model)
    for signed in products:
        if not check_product(fm, signed):
            raise AssertionError("sampler produced an invalid product")
    write_pdt(output, fm, products)
    return len(products)


def main(argv=None):
    """Command line with PLEDGE's own flags, so that the ``java`` stand-in of featurenet_b200/compat/bin can forward
    the reference's argument list unchanged:  generate_products -fm X.xml -nbProds N -o OUT.pdt [-timeAllowedMS T]"""
    import sys
    argv = list(sys.argv[1:] if argv is None else argv)
    if "-jar" in argv:                       # called as `java -jar <path to PLEDGE.jar> ...`
        at = argv.index("-jar")
        del argv[at: at + 2]
    if not argv or argv[0] != "generate_products":
        raise SystemExit("usage: generate_products -fm <model.xml> -nbProds <n> -o <out.pdt> [-timeAllowedMS <ms>] [-seed <s>]")
    opts, i = {}, 1
    while i + 1 < len(argv):
        opts[argv[i]] = argv[i + 1]
        i += 2
    for needed in ("-fm", "-nbProds", "-o"):
        if needed not in opts:
            raise SystemExit("missing %s" % needed)
    n = generate_products(opts["-fm"], int(opts["-nbProds"]), opts["-o"],
                          seed=int(opts["-seed"]) if "-seed" in opts else None)
    print("%d products written to %s" % (n, opts["-o"]))
    return 0


if __name__ == "__main__":
    raise SystemExit(main())
