"""cuBLAS DGEMM throughput (the FP64 'tensor roofline' denominator), burst and sustained."""
import json
import time

import torch

n = 8192
a = torch.randn(n, n, dtype=torch.float64, device="cuda")
b = torch.randn(n, n, dtype=torch.float64, device="cuda")
for _ in range(3):
    torch.matmul(a, b)
torch.cuda.synchronize()
best = 1e9
for _ in range(5):
    e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
    e0.record(); torch.matmul(a, b); e1.record(); torch.cuda.synchronize()
    best = min(best, e0.elapsed_time(e1))
burst = 2 * n ** 3 / best * 1e-9
e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
t = time.time(); k = 0
e0.record()
while time.time() - t < 3.0:
    for _ in range(5):
        torch.matmul(a, b); k += 1
    torch.cuda.synchronize()
e1.record(); torch.cuda.synchronize()
sus = 2 * n ** 3 * k / e0.elapsed_time(e1) * 1e-9
print(json.dumps({"probe": "cublas_dgemm_8192", "burst_tflops": round(burst, 2), "sustained_tflops": round(sus, 2)}))
