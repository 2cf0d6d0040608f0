"""Accuracy of the 3xTF32 path of fp32 feature maps against the FP64-DMMA path and the fp64 oracle."""
import os, sys, subprocess
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from libskylark_b200 import lib, sketch
from oracle import oracle as o

if len(sys.argv) > 1:      # child: compute and save
    m, n, s = 20000, 784, 16384
    X = torch.rand(m, n, dtype=torch.float32, device="cuda", generator=torch.Generator("cuda").manual_seed(0))
    S = sketch.GaussianRFT(n, s, 10.0, context=lib.Context(123834))
    np.save(sys.argv[1], S.apply(X, None, "rowwise").cpu().numpy())
    sys.exit(0)
m, n, s = 20000, 784, 16384
env = dict(os.environ)
subprocess.check_call([sys.executable, __file__, "/tmp/tf32.npy"], env=env)
env["SKB_DEBUG_NO_TF32"] = "1"
subprocess.check_call([sys.executable, __file__, "/tmp/dmma.npy"], env=env)
a, b = np.load("/tmp/tf32.npy").astype(np.float64), np.load("/tmp/dmma.npy").astype(np.float64)
X = torch.rand(m, n, dtype=torch.float32, device="cuda", generator=torch.Generator("cuda").manual_seed(0)).cpu().numpy()
O = o.GaussianRFT(n, s, 10.0, o.Context(123834))
ref = O.apply(X[:512].astype(np.float64), "rowwise")
fro = lambda x, y: np.linalg.norm(x - y) / np.linalg.norm(y)
print("3xTF32 vs fp64 oracle  (512 points): %.3e" % fro(a[:512], ref))
print("DMMA   vs fp64 oracle  (512 points): %.3e" % fro(b[:512], ref))
print("3xTF32 vs DMMA path (20000 points): %.3e   max abs %.3e" % (fro(a, b), np.abs(a - b).max()))
