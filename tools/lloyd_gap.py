"""Where a 20-iteration fit's wall time goes (cfg5 shapes): device time of the STORE pass and of the REUSE passes enqueued back to
back, against the wall time of ml.clustering.lloyd (host loop with its state read-backs).  python tools/lloyd_gap.py [n d]"""
import os, sys, time, json
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
import hkm_b200
from hkm_b200 import datagen, _native as N, _runtime as RT
from hkm_b200.ml import clustering as CL

n = int(sys.argv[1]) if len(sys.argv) > 1 else 5_000_000
d = int(sys.argv[2]) if len(sys.argv) > 2 else 24
dev = RT.device()
X = datagen.hospital_matrix_device(n, d, 8, 20250331, dev)
for k in (4, 8, 12, 16, 24, 32):
    C0 = X[:k].cpu().numpy()
    for rep in range(2):
        torch.cuda.synchronize(); t0 = time.perf_counter()
        C, it, cost = CL.lloyd(X, C0, None, 20, 0.0)
        torch.cuda.synchronize(); wall = (time.perf_counter() - t0) * 1e3
    C = torch.from_numpy(C0).to(dev); P = k * d + k + 1
    partial = torch.zeros(P, dtype=torch.float64, device=dev); scratch = N.new_scratch(d, k, dev); state = N.new_state(dev)
    xs = torch.zeros(1, dtype=torch.float64, device=dev)
    ev = [torch.cuda.Event(enable_timing=True) for _ in range(3)]
    for rep in range(2):
        C.copy_(torch.from_numpy(C0)); state.zero_()
        torch.cuda.synchronize(); ev[0].record()
        N.lloyd_step(X, C, partial, 0.0, state, scratch, xnorm2_sum=xs, xnorm2_mode=1)
        ev[1].record()
        for _ in range(19):
            N.lloyd_step(X, C, partial, 0.0, state, scratch, xnorm2_sum=xs, xnorm2_mode=2)
        ev[2].record(); torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(200):
        N.lloyd_step(X[:256], C, partial, 0.0, state, scratch, xnorm2_sum=xs, xnorm2_mode=0)
    host_us = (time.perf_counter() - t0) / 200 * 1e6
    torch.cuda.synchronize()
    print(json.dumps(dict(n=n, d=d, k=k, iters=int(it), wall_ms=round(wall, 3), store_pass_ms=round(ev[0].elapsed_time(ev[1]), 3),
                          reuse_19_ms=round(ev[1].elapsed_time(ev[2]), 3), reuse_per_iter_ms=round(ev[1].elapsed_time(ev[2]) / 19, 4),
                          hbm_ms=round(n * d * 8 / 6554.6e6, 4), host_us_per_call=round(host_us, 1))), flush=True)
