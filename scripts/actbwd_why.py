import sys, os, contextlib, io, copy, collections
sys.path.insert(0, '/root/repo')
import featurenet_b200; featurenet_b200.install_compat()
from products_tree import ProductSet
from model.keras_model import KerasFeatureModel
from featurenet_b200.optimizers import Adam
from featurenet_b200 import lib as fl
with contextlib.redirect_stdout(io.StringIO()):
    formatted = ProductSet("/root/repo/featurenet_b200/data/nas_5_5_10_pop32_seed7.pdt").format_products()
why = collections.Counter()
for idx, (product, _) in enumerate(formatted):
    with contextlib.redirect_stdout(io.StringIO()):
        net = KerasFeatureModel.parse_feature_model(copy.deepcopy(product)).build((32, 32, 3), 10)
    if not net: continue
    net.compile(loss="categorical_crossentropy", optimizer=Adam(1e-3))
    eng = net._need_engine()
    eng._ensure_capacity(128, 128)
    consumers = collections.defaultdict(list)
    for j in range(1, len(eng.nodes)):
        if eng.nodes[j].kind == "flatten": continue
        for s in eng.src[j]:
            consumers[eng._resolve(s, True)].append(j)
    need = eng._needs_grad()
    for r, node in enumerate(eng.nodes):
        if node.kind in ("conv2d", "sepconv2d", "dense", "dwconv2d") and fl.ACT[node.cfg["activation"]] in (1,2,3) and r != len(eng.nodes)-1 and need[r]:
            users = consumers.get(r, [])
            kinds = tuple(sorted(eng.nodes[u].kind for u in users))
            if len(users) == 1 and eng.nodes[users[0]].kind in ("conv2d", "dense") and len(eng.src[users[0]]) == 1:
                tc = eng._dgrad_on_tensor_core(users[0], 128)
                why["single conv/dense consumer, tensor-core dgrad" if tc else "FUSED"] += 1
            else:
                why["consumers %s" % (kinds,)] += 1
    net.release()
for k, v in why.most_common(): print(v, k)
