#!/bin/bash
# Runs on the GPU box: FP64 microbenchmarks + cuBLAS DGEMM through torch (library reference for the FP64 roofline).
mkdir -p gpurun_out
./tools/microbench_fp64 | tee gpurun_out/microbench_fp64.txt
python - <<'PY' | tee -a gpurun_out/microbench_fp64.txt
import torch, time
a = torch.randn(8192, 8192, dtype=torch.float64, device="cuda"); b = torch.randn_like(a)
for _ in range(2): (a @ b)
best = 1e9
for _ in range(5):
    e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
    e0.record(); c = a @ b; e1.record(); torch.cuda.synchronize(); best = min(best, e0.elapsed_time(e1))
print(f"cuBLAS DGEMM 8192^3: {2*8192**3/best/1e9:.2f} TFLOP/s (best of 5)")
a = torch.randn(4096, 4096, dtype=torch.float64, device="cuda"); a = a @ a.T + 4096*torch.eye(4096, dtype=torch.float64, device="cuda")
torch.linalg.cholesky(a); torch.cuda.synchronize()
e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
e0.record(); torch.linalg.cholesky(a); e1.record(); torch.cuda.synchronize()
print(f"cuSOLVER potrf n=4096: {e0.elapsed_time(e1):.3f} ms")
PY
