"""Where does the host time of a cfg3 shard step go?  perf_counter around the C call and around the group average."""
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
tc = tg = 0.0; n = 20
t00 = time.perf_counter()
for i in range(n):
    t0 = time.perf_counter(); r = gibbs.run_batch(pb, cfg.nsamp, handle=h, seed=10 + i); t1 = time.perf_counter()
    avg, cnt = ddist.group_average(r.spectrum, gids, 32, 64, weights=r.success); torch.cuda.synchronize(); t2 = time.perf_counter()
    tc += t1 - t0; tg += t2 - t1
print(f"per step: total {(time.perf_counter()-t00)/n*1e3:.2f} ms, run_batch {tc/n*1e3:.2f} ms, group_average+sync {tg/n*1e3:.2f} ms; kernels {r.timing}")
import cProfile, pstats
pr = cProfile.Profile(); pr.enable()
for i in range(10):
    r = gibbs.run_batch(pb, cfg.nsamp, handle=h, seed=50 + i)
pr.disable(); pstats.Stats(pr).sort_stats("cumulative").print_stats(14)
