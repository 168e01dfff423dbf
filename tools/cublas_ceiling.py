#!/usr/bin/env python
"""What does the vendor GEMM reach on the LTV MLP's shapes on this GPU?  (A ceiling for k_mlp_gemm's layers;
library code, not part of the product path.)"""
import json, torch
dev = torch.device("cuda", 0)
out = {}
for name, (m, k, n) in {"layer1 256->1024": (1 << 20, 256, 1024), "layer2/3 1024->1024": (1 << 20, 1024, 1024)}.items():
    a = torch.randn(m, k, device=dev, dtype=torch.bfloat16)
    w = torch.randn(n, k, device=dev, dtype=torch.bfloat16)
    for _ in range(5):
        c = a @ w.t()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    reps = 30
    for _ in range(reps):
        c = a @ w.t()
    e1.record(); torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / reps
    out[name] = {"m": m, "k": k, "n": n, "ms": ms, "tflops": 2.0 * m * k * n / (ms * 1e-3) / 1e12}
print(json.dumps(out))
