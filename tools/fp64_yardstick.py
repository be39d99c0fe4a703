"""cuBLAS DGEMM + HBM copy yardsticks (library calls: yardstick only, never on the product path)."""
import json, torch, time
torch.backends.cuda.matmul.allow_tf32 = False
dev = torch.device("cuda:0")
res = {"gpu": torch.cuda.get_device_name(0)}
for n in (4096, 8192):
    a = torch.randn(n, n, device=dev, dtype=torch.float64); b = torch.randn(n, n, device=dev, dtype=torch.float64)
    for _ in range(2): (a @ b)
    best = 1e9
    for _ in range(5):
        e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
        e0.record(); c = a @ b; e1.record(); torch.cuda.synchronize()
        best = min(best, e0.elapsed_time(e1))
    res[f"dgemm_{n}_tflops_burst"] = 2 * n**3 / best * 1e-9
n = 8192
torch.cuda.synchronize(); e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
e0.record(); k = 0
t = time.time()
while time.time() - t < 3.0:
    c = a @ b; k += 1
e1.record(); torch.cuda.synchronize()
res["dgemm_8192_tflops_sustained"] = 2 * n**3 * k / e0.elapsed_time(e1) * 1e-9
# TN-shaped like the eris GEMM: (32768 x 1536) x (1536 x 8192)
A = torch.randn(32768, 1536, device=dev, dtype=torch.float64)
best = 1e9
for _ in range(4):
    e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
    e0.record(); c = A @ A[:8192].T; e1.record(); torch.cuda.synchronize(); best = min(best, e0.elapsed_time(e1))
res["dgemm_32768x8192x1536_tflops"] = 2 * 32768 * 8192 * 1536 / best * 1e-9
x = torch.empty(1 << 30, device=dev, dtype=torch.float64); y = torch.empty_like(x)
best = 1e9
for _ in range(6):
    e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
    e0.record(); y.copy_(x); e1.record(); torch.cuda.synchronize(); best = min(best, e0.elapsed_time(e1))
res["hbm_copy_gbs"] = 2 * x.numel() * 8 / best * 1e-6
print(json.dumps(res))
