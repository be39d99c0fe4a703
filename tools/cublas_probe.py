import torch
a = torch.randn(8192, 1536, device="cuda", dtype=torch.float64)
for _ in range(3):
    c = a @ a.T
torch.cuda.synchronize()
