import sys, statistics, torch
sys.path.insert(0, '.')
from cuda_mpi_openmp_b200 import ops
from cuda_mpi_openmp_b200.utils import cuda_time_ms
dev = torch.device('cuda:0')
for h, w in [(16384, 16384), (4096, 65536), (65536, 4096), (8192, 32768), (32768, 8192), (16384, 65536)]:
    img = torch.randint(-2**31, 2**31 - 1, (h, w), dtype=torch.int32, device=dev)
    out = torch.empty_like(img)
    t = cuda_time_ms(lambda: ops.roberts(img, out, variant='tma'), iters=10, warmup=3)
    ms = statistics.median(t)
    print(f"{w}x{h}: {ms:.4f} ms  {8*h*w/ms/1e6:.0f} GB/s")
    del img, out
