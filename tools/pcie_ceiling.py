"""Raw ceiling of the end-to-end leg: every rank copies page-locked host memory to its GPU and
back at the same time (one cudaMemcpyAsync per direction per chunk, two streams), all ranks concurrently.
Run under torchrun at N = 1, 2, 4, 8; rank 0 prints one JSON line: GB/s each way per GPU and in total, beside the
rate b2c_convolve_batch_host reaches on the headline shape (80 000 B each way per convolution)."""
import json, os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch, torch.distributed as dist

rank, world, local = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1)), int(os.environ.get("LOCAL_RANK", 0))
torch.cuda.set_device(local)
if world > 1:
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))

def barrier():
    if world > 1: dist.barrier()
    torch.cuda.synchronize()

def allmax(x):
    t = torch.tensor([x], dtype=torch.float64, device="cuda")
    if world > 1: dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())

MB = 1024
n = MB << 20
h_in, h_out = torch.empty(n, dtype=torch.uint8).pin_memory(), torch.empty(n, dtype=torch.uint8).pin_memory()
d_in, d_out = torch.empty(n, dtype=torch.uint8, device="cuda"), torch.empty(n, dtype=torch.uint8, device="cuda")
s1, s2 = torch.cuda.Stream(), torch.cuda.Stream()
res = {}
for name, do_in, do_out in (("h2d_only", True, False), ("d2h_only", False, True), ("both", True, True)):
    best = None
    for it in range(4):
        barrier()
        t0 = time.perf_counter()
        if do_in:
            with torch.cuda.stream(s1): d_in.copy_(h_in, non_blocking=True)
        if do_out:
            with torch.cuda.stream(s2): h_out.copy_(d_out, non_blocking=True)
        torch.cuda.synchronize()
        dt = allmax(time.perf_counter() - t0)
        if it: best = dt if best is None else min(best, dt)
    res[name] = n / best / 1e9

# the pipeline itself on the same box, same ranks
import fftconvolution_b200 as b2c
b2c.lib().b2c_set_device(local)
N_IMG = 32768
plan = b2c.Plan(b2c.LINEAR_SAME, 100, 100, 15, 15)
hs = torch.rand((N_IMG, 100, 100), dtype=torch.float64).pin_memory()
hd = torch.empty((N_IMG, 100, 100), dtype=torch.float64).pin_memory()
hk = torch.rand((15, 15), dtype=torch.float64).pin_memory()
plan.convolve_batch_host_ptr(hs.data_ptr(), hk.data_ptr(), 1, hd.data_ptr(), N_IMG)
barrier()
t0 = time.perf_counter()
for _ in range(3): plan.convolve_batch_host_ptr(hs.data_ptr(), hk.data_ptr(), 1, hd.data_ptr(), N_IMG)
torch.cuda.synchronize()
dt = allmax(time.perf_counter() - t0)
pipe = 3 * N_IMG / dt
if rank == 0:
    out = {"n_gpus": world, "copy_mib_per_direction": MB, "gbs_per_gpu": res, "gbs_total_both": res["both"] * world,
           "ceiling_conv_s_total": res["both"] * world * 1e9 / 80000.0, "pipeline_conv_s_total": pipe * world,
           "pipeline_frac_of_ceiling": pipe * world / (res["both"] * world * 1e9 / 80000.0),
           "note": "gbs_per_gpu = each direction, slowest rank; 'both' = H2D and D2H running concurrently"}
    print(json.dumps(out), flush=True)
if world > 1: dist.destroy_process_group()
