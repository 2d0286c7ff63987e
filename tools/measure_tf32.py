"""Dense TF32 peak with the protocol of MEASURED_PEAKS.json's bf16 number (torch.matmul 8192^3, allow_tf32, best of 10 and a
4 s sustained run) -- recorded in BASELINE.md; bench.py keeps 0.5 x the measured bf16 peak as the roofline denominator."""
import json, time, torch
torch.backends.cuda.matmul.allow_tf32 = True
n = 8192
a = torch.randn(n, n, device="cuda"); b = torch.randn(n, n, device="cuda")
for _ in range(3):
    a @ b
torch.cuda.synchronize()
best = 0.0
for _ in range(10):
    s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    s.record(); a @ b; e.record(); e.synchronize()
    best = max(best, 2 * n ** 3 / (s.elapsed_time(e) * 1e-3) / 1e12)
t0 = time.time(); k = 0
s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
s.record()
while time.time() - t0 < 4.0:
    for _ in range(20):
        a @ b
    k += 20
    torch.cuda.synchronize()
e.record(); e.synchronize()
print(json.dumps({"tf32_tflops_burst": best, "tf32_tflops_sustained": 2 * n ** 3 * k / (s.elapsed_time(e) * 1e-3) / 1e12}))
