import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch, fftconvolution_b200 as b2c
torch.cuda.set_device(0)
n = 16384
p = b2c.Plan(2, 1, 4096, 1, 255); p.set_path(b2c.PATH_FFT)
src = torch.rand((n, 1, 4096), dtype=torch.float64, device="cuda"); ker = torch.rand((1, 255), dtype=torch.float64, device="cuda")
dst = torch.empty_like(src)
for _ in range(3): p.convolve_batch_device(src.data_ptr(), ker.data_ptr(), 1, dst.data_ptr(), n, 0)
torch.cuda.synchronize()
