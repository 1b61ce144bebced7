import sys; sys.path.insert(0,'/root/repo')
import torch, numpy as np, fftconvolution_b200 as b2c
torch.cuda.set_device(0); st=torch.cuda.current_stream().cuda_stream
n=int(sys.argv[1]) if len(sys.argv)>1 else 8192
plan=b2c.Plan(b2c.CIRCULAR_FULL,100,100,100,100)
src=torch.empty((n,100,100),dtype=torch.float64,device='cuda'); ker=torch.empty_like(src); dst=torch.empty((n,199,199),dtype=torch.float64,device='cuda')
b2c.fill_uniform_device(src.data_ptr(),src.numel(),3); b2c.fill_uniform_device(ker.data_ptr(),ker.numel(),4)
for _ in range(2): plan.convolve_batch_device(src.data_ptr(),ker.data_ptr(),0,dst.data_ptr(),n,st)
torch.cuda.synchronize()
e0,e1=torch.cuda.Event(enable_timing=True),torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(3): plan.convolve_batch_device(src.data_ptr(),ker.data_ptr(),0,dst.data_ptr(),n,st)
e1.record(); torch.cuda.synchronize()
ms=e0.elapsed_time(e1)/3
print('config3: %.3f ms  %.3f M conv/s  %.1f%% of HBM roof'%(ms, n/ms/1e3, 100*n/ms*1e3*476808/6540.2e9))
