"""Copy bandwidth of THIS GPU, measured the way MEASURED_PEAKS.json states (torch b.copy_(a), read+write bytes)."""
import torch
n = 1 << 30
a = torch.empty(n, dtype=torch.bfloat16, device="cuda").normal_()
b = torch.empty_like(a)
best = 1e9
for _ in range(12):
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(); b.copy_(a); e1.record(); torch.cuda.synchronize()
    best = min(best, e0.elapsed_time(e1))
print(f"copy: {2 * n * 2 / best / 1e6:.1f} GB/s (best of 12, {best:.4f} ms)")
