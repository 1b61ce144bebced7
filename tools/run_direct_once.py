import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch, fftconvolution_b200 as b2c
torch.cuda.set_device(0)
hk = int(sys.argv[1]) if len(sys.argv) > 1 else 3
hs = int(sys.argv[2]) if len(sys.argv) > 2 else 100
n = 16384
p = b2c.Plan(2, hs, hs, hk, hk); p.set_path(b2c.PATH_DIRECT)
src = torch.rand((n, hs, hs), dtype=torch.float64, device="cuda"); ker = torch.rand((hk, hk), dtype=torch.float64, device="cuda")
dst = torch.empty_like(src)
for _ in range(3): p.convolve_batch_device(src.data_ptr(), ker.data_ptr(), 1, dst.data_ptr(), n, 0)
torch.cuda.synchronize()
