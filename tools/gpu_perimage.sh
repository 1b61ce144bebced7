cd /root/repo
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; tail -4 gpurun_out/pytest_gpu.log
timeout 300 python tools/bench_perimage.py 2>&1 | tee gpurun_out/perimage.txt
timeout 300 python tools/quick_bench.py 65536 20 2>&1 | tee gpurun_out/ab.txt
