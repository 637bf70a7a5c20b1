import sys
import torch
sys.path.insert(0, ".")
from draco_b200 import gibbs
N = int(sys.argv[1]) if len(sys.argv) > 1 else 2048
x = torch.randn((1 << 26) // N, N, 2, dtype=torch.float64, device="cuda")
for _ in range(2):
    gibbs.fft_rows(x, "nuttall", shift=True)
torch.cuda.synchronize()
