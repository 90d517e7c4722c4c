import importlib, os, sys, ctypes
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
N = importlib.import_module("clustermachinelearningforhospitalnetworks-apache-spark_b200._native")
v = int(sys.argv[1]); n, d, k = int(sys.argv[2]), int(sys.argv[3]), int(sys.argv[4])
mode = int(sys.argv[5]) if len(sys.argv) > 5 else -1
iters = int(sys.argv[6]) if len(sys.argv) > 6 else 20
if os.environ.get("HKM_WG_GROUPS"):
    N.load().hkm_debug_wg_groups(ctypes.c_int(int(os.environ["HKM_WG_GROUPS"])))
X = torch.randn(n, d, dtype=torch.float64, device="cuda"); C = X[:k].clone()
partial = torch.empty(k * d + k + 1, dtype=torch.float64, device="cuda"); scratch = N.new_scratch(d, k, "cuda")
xs = torch.zeros(1, dtype=torch.float64, device="cuda")
N.set_lloyd_variant(v)
i = -1
try:
    for i in range(iters):
        m = mode if mode >= 0 else (1 if i < 3 else 2)
        if m == 2 and i == 0:
            xs.fill_(float((X * X).sum()))
        N.assign_reduce(X, C, partial, scratch=scratch, xnorm2_sum=xs, xnorm2_mode=m)
        torch.cuda.synchronize()
    print("ok", v, n, d, k, "mode", mode, partial[-1].item())
except Exception as ex:
    print("FAIL at launch", i, "v", v, n, d, k, "mode", mode, str(ex)[:80].replace("\n", " "))
