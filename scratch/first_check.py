import sys, time, numpy as np
sys.path.insert(0, "/root/repo")
import oracle
from draco_b200 import gibbs

def rel(a, b):
    return np.max(np.abs(a - b)) / max(np.max(np.abs(b)), 1e-300)

rng = np.random.default_rng(0)
# 1. fourier matrices
N = 32; fsel = np.array([0, 1, 3, 7, 16])
print("r2c", rel(gibbs.fourier_matrix_r2c(N, fsel), oracle.dft_r2c(N, fsel)))
print("c2r", rel(gibbs.fourier_matrix_c2r(N, fsel), oracle.dft_c2r(N, fsel)))
print("c2c", rel(gibbs.fourier_matrix_c2c(N, fsel), oracle.dft_c2c(N, fsel)))
print("cpx", rel(gibbs.fourier_matrix(N, fsel), oracle.dft_complex(N, fsel)))

def case(N, M, T, cplx, window, niter, seed, nyq=False):
    r = np.random.default_rng(seed)
    nch = N if cplx else N // 2 + 1
    fsel = np.sort(r.choice(nch - (0 if nyq or cplx else 1), size=M, replace=False))
    data = r.standard_normal((T, M)) + 1j * r.standard_normal((T, M))
    Ni = r.uniform(0.5, 2.0, M) * 100
    S0 = np.full(N, 10.0)
    return data, Ni, fsel, S0

for (N, M, T, cplx, window, niter) in [(64, 29, 8, False, "nuttall", 6), (64, 33, 8, False, None, 4), (256, 102, 32, False, "nuttall", 20),
                                       (64, 50, 8, True, "nuttall", 5), (100, 40, 5, False, "hann", 4), (300, 150, 12, False, "blackman_harris", 3)]:
    data, Ni, fsel, S0 = case(N, M, T, cplx, window, niter, 1, nyq=(window is None))
    z, fh, ff = oracle.whiten_inputs(data, N, Ni, fsel, window, cplx)
    dz, dfh, dff = gibbs.compute_delay_spectrum_inputs(data, N, Ni, fsel, window, cplx)
    print(f"N={N} M={M} T={T} cplx={cplx} win={window}: setup rel err z={rel(dz,z):.2e} FTNih={rel(dfh,fh):.2e} FTNiF={rel(dff,ff):.2e}")
    wf = oracle.wiener_filter(S0, data, N, Ni, window, fsel, cplx)
    dwf = gibbs.delay_spectrum_wiener_filter(S0, data, N, Ni, window, fsel, cplx)
    print(f"    wiener rel err {rel(dwf, wf):.2e}")
    rec = oracle.RecordingDraws(np.random.default_rng(7))
    so, oko = oracle.gibbs_chain(data, N, Ni, S0, window, fsel, niter, rec, cplx)
    sd, okd = gibbs.delay_power_spectrum_gibbs(data, N, Ni, S0, window, fsel, niter, np.random.default_rng(7), cplx)
    so, sd = np.array(so), np.array(sd)
    print(f"    gibbs ok={oko}/{okd} n={len(so)}/{len(sd)} max rel dev per-sample {np.max(np.abs(sd/so-1)):.2e}")

# timing probe
for (B, N, T, niter) in [(148, 2048, 64, 2)]:
    M = N // 2
    r = np.random.default_rng(5)
    vis = (r.standard_normal((B, T, M)) + 1j * r.standard_normal((B, T, M)))
    w = r.uniform(0.5, 2, (B, M)) * 1e2
    pb = gibbs.pack_uniform(vis, w, np.arange(M), N, np.full((B, N), 10.0)).to_device(0)
    for rep in range(2):
        t0 = time.time()
        res = gibbs.run_batch(pb, niter, seed=1)
        tm = res.timing
        print(f"B={B} N={N} T={T} niter={niter}: wall {time.time()-t0:.3f}s sampler {tm['ms_sampler']:.1f} ms setup {tm['ms_setup']:.1f} ms "
              f"-> {tm['flop_sampler']/tm['ms_sampler']/1e9:.2f} TFLOP/s, {B*niter/tm['ms_sampler']*1e3:.1f} baseline-iter/s; ok={int(res.success.sum())}")
