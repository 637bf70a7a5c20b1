import torch, time
torch.backends.cuda.matmul.allow_tf32 = False
dev = "cuda:0"
print(torch.cuda.get_device_name(0))
for n in (4096, 8192):
    a = torch.randn(n, n, dtype=torch.float64, device=dev)
    b = torch.randn(n, n, dtype=torch.float64, device=dev)
    for _ in range(2): c = a @ b
    torch.cuda.synchronize()
    best = 1e9
    for _ in range(5):
        e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
        e0.record(); c = a @ b; e1.record(); torch.cuda.synchronize()
        best = min(best, e0.elapsed_time(e1))
    print(f"cuBLAS DGEMM {n}^3: {best:.3f} ms  {2*n**3/best/1e9:.2f} TFLOP/s")
# sustained
n = 8192
t0 = time.time(); cnt = 0
e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
e0.record()
while cnt < 40:
    c = a @ b; cnt += 1
e1.record(); torch.cuda.synchronize()
print(f"cuBLAS DGEMM 8192^3 sustained x{cnt}: {2*n**3*cnt/e0.elapsed_time(e1)/1e9:.2f} TFLOP/s")
# batched cholesky via cuSOLVER/MAGMA for reference (library baseline)
for (bsz, n) in ((64, 2048), (256, 256)):
    m = torch.randn(bsz, n, n, dtype=torch.float64, device=dev)
    spd = m @ m.transpose(1, 2) + n * torch.eye(n, dtype=torch.float64, device=dev)
    torch.linalg.cholesky(spd); torch.cuda.synchronize()
    e0.record(); L = torch.linalg.cholesky(spd); e1.record(); torch.cuda.synchronize()
    ms = e0.elapsed_time(e1)
    print(f"torch.linalg.cholesky batch={bsz} n={n}: {ms:.2f} ms -> {bsz*n**3/3/ms/1e9:.2f} TFLOP/s, {bsz/ms*1e3:.1f} fact/s")
