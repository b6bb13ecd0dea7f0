"""tf32x3 / auto products of one Dense shape against float64 (development aid): python scripts/x3_check.py M K N [mode]"""
import ctypes as C, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np, torch
from featurenet_b200 import lib as fl
L = fl.get()
m, k, n = (int(v) for v in sys.argv[1:4])
mode = fl.MATH[sys.argv[4] if len(sys.argv) > 4 else "tf32x3"]
rng = np.random.default_rng(0)
x, w, dz = rng.standard_normal((m, k)), rng.standard_normal((k, n)) / np.sqrt(k), rng.standard_normal((m, n))
dev = lambda a: torch.from_numpy(a.astype(np.float32)).cuda()
ptr = lambda t: C.c_void_p(t.data_ptr())
xd, wd, dzd = dev(x), dev(w), dev(dz)
x, w, dz = (t.cpu().numpy().astype(np.float64) for t in (xd, wd, dzd))
ws = torch.empty(max(int(L.fnet_dense_workspace_bytes(m, k, n)), 256), dtype=torch.uint8, device="cuda")
y, dx, dw, db = torch.empty(m, n, device="cuda"), torch.empty(m, k, device="cuda"), torch.empty(k, n, device="cuda"), torch.empty(n, device="cuda")
L.fnet_dense_fwd(m, k, n, 0, mode, ptr(xd), ptr(wd), None, ptr(y), ptr(ws), ws.numel(), None)
L.fnet_dense_bwd_data(m, k, n, mode, ptr(dzd), ptr(wd), ptr(dx), 0, ptr(ws), ws.numel(), None)
L.fnet_dense_bwd_filter(m, k, n, mode, ptr(xd), ptr(dzd), ptr(dw), ptr(db), ptr(ws), ws.numel(), None)
torch.cuda.synchronize()
rel = lambda a, b: np.abs(a.cpu().numpy() - b).max() / np.abs(b).max()
print("fwd %.2e  dgrad %.2e  wgrad %.2e  dbias %.2e" % (rel(y, x @ w), rel(dx, dz @ w.T), rel(dw, x.T @ dz), rel(db, dz.sum(0))))
