import sys, time
import numpy as np, torch
sys.path.insert(0, ".")
import bench
from draco_b200 import gibbs, synth, dist as ddist
cfg = synth.CONFIGS["cfg3"]; nbase = 2048
pb = bench.make_inputs(cfg, nbase, 0).to_device(0)
h = gibbs.Handle(0, stream=torch.cuda.current_stream().cuda_stream)
y = torch.randn(2048, 2048, dtype=torch.float64, device="cuda")
gids = np.arange(nbase)
for i in range(3):
    r = gibbs.run_batch(pb, cfg.nsamp, handle=h, seed=i)
torch.cuda.synchronize()
def tm(f):
    t0 = time.perf_counter(); f(); torch.cuda.synchronize(); return (time.perf_counter() - t0) * 1e3
rows = []
st0 = torch.cuda.memory_stats()
for i in range(8):
    r = gibbs.run_batch(pb, cfg.nsamp, handle=h, seed=10 + i); torch.cuda.synchronize()
    s = r.spectrum
    rows.append([tm(lambda: y.sum()), tm(lambda: s.sum()), tm(lambda: s.sum()), tm(lambda: s * 2.0), tm(lambda: torch.zeros((32, 2049), dtype=torch.float64, device="cuda")),
                 tm(lambda: ddist.group_average(s, gids, 32, 64, weights=r.success)), tm(lambda: ddist.group_average(s, gids, 32, 64, weights=r.success))])
st1 = torch.cuda.memory_stats()
print("y.sum, s.sum, s.sum, s*2, zeros, group_average, group_average [ms, median of 8]:", np.round(np.median(np.array(rows), axis=0), 3))
print("first rows:", np.round(np.array(rows[:3]), 3))
for k in ("num_device_alloc", "num_device_free", "num_alloc_retries", "num_sync_all_streams"):
    print(k, st0.get(k), st1.get(k))
print("reserved GB", torch.cuda.memory_reserved() / 1e9, "free/total", [x / 1e9 for x in torch.cuda.mem_get_info()])
