import sys, time
import numpy as np, torch
sys.path.insert(0, ".")
B, N, ng, gs = 2048, 2048, 32, 64
local = torch.randn(B, N, dtype=torch.float64, device="cuda"); ok = torch.ones(B, dtype=torch.uint8, device="cuda")
gids = np.arange(B)
def T(label, f, n=20):
    for _ in range(3): f()
    torch.cuda.synchronize(); t0 = time.perf_counter()
    for _ in range(n): r = f()
    torch.cuda.synchronize(); print(f"{label:40s} {(time.perf_counter()-t0)/n*1e3:8.3f} ms"); return r
gid = T("as_tensor(gid)", lambda: torch.as_tensor(np.asarray(gids) // gs, device="cuda", dtype=torch.long))
w = T("weights.to", lambda: ok.to(torch.float64))
acc = torch.zeros((ng, N + 1), dtype=torch.float64, device="cuda")
T("zeros", lambda: torch.zeros((ng, N + 1), dtype=torch.float64, device="cuda"))
T("local*w", lambda: local * w[:, None])
lw = local * w[:, None]
T("index_add_ strided", lambda: acc[:, :-1].index_add_(0, gid, lw))
acc2 = torch.zeros((ng, N), dtype=torch.float64, device="cuda")
T("index_add_ contiguous", lambda: acc2.index_add_(0, gid, lw))
T("index_add_ count", lambda: acc[:, -1].index_add_(0, gid, w))
G = torch.zeros((ng, B), dtype=torch.float64, device="cuda")
def onehot():
    G = torch.zeros((ng, B), dtype=torch.float64, device="cuda"); G[gid, torch.arange(B, device="cuda")] = w; return G
G = T("one-hot build", onehot)
T("G @ local", lambda: G @ local)
from draco_b200 import dist as ddist
T("group_average (current)", lambda: ddist.group_average(local, gids, ng, gs, weights=ok))
