import sys, time
import numpy as np, torch
sys.path.insert(0, ".")
import bench
from draco_b200 import gibbs, synth
def probe(cfgname, nbase, nsamp):
    cfg = synth.CONFIGS[cfgname]
    pb = bench.make_inputs(cfg, nbase, 0).to_device(0)
    h = gibbs.Handle(0, stream=torch.cuda.current_stream().cuda_stream)
    y = torch.randn(2048, 2048, dtype=torch.float64, device="cuda")
    for i in range(3):
        r = gibbs.run_batch(pb, nsamp, handle=h, seed=i)
    torch.cuda.synchronize()
    rows = []
    for i in range(6):
        r = gibbs.run_batch(pb, nsamp, handle=h, seed=10 + i); torch.cuda.synchronize()
        ts = []
        for k in range(6):
            t0 = time.perf_counter(); x = y.sum(); torch.cuda.synchronize(); ts.append((time.perf_counter() - t0) * 1e3)
        rows.append(ts)
    print(cfgname, nbase, "ms of 6 consecutive y.sum()+sync after run_batch (median over 6 runs):", np.round(np.median(np.array(rows), axis=0), 3))
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    r = gibbs.run_batch(pb, nsamp, handle=h, seed=99); torch.cuda.synchronize()
    e0.record(); x = y.sum(); e1.record(); torch.cuda.synchronize(); print("  device time of the first y.sum():", round(e0.elapsed_time(e1), 3), "ms")
    del h
probe("cfg3", 2048, 20)
probe("cfg3", 256, 20)
probe("cfg2", 148, 2)
