import sys, time
import numpy as np, torch
sys.path.insert(0, ".")
import bench
from draco_b200 import gibbs, synth, dist as ddist
cfg = synth.CONFIGS["cfg3"]; nbase = 2048
pb = bench.make_inputs(cfg, nbase, 0).to_device(0)
h = gibbs.Handle(0, stream=torch.cuda.current_stream().cuda_stream)
gids = np.arange(nbase)
for i in range(3):
    r = gibbs.run_batch(pb, cfg.nsamp, handle=h, seed=i)
torch.cuda.synchronize()
acc = np.zeros(6); n = 20
for i in range(n):
    t = [time.perf_counter()]
    r = gibbs.run_batch(pb, cfg.nsamp, handle=h, seed=10 + i); t.append(time.perf_counter())
    torch.cuda.synchronize(); t.append(time.perf_counter())
    s = r.spectrum; ok = r.success; t.append(time.perf_counter())
    x = s.sum(); torch.cuda.synchronize(); t.append(time.perf_counter())
    avg, cnt = ddist.group_average(s, gids, 32, 64, weights=ok); torch.cuda.synchronize(); t.append(time.perf_counter())
    del r, s, ok, avg, cnt, x; torch.cuda.synchronize(); t.append(time.perf_counter())
    acc += np.diff(t)
print("run_batch, sync, attr, first torch op, group_average, del:", np.round(acc / n * 1e3, 3))
