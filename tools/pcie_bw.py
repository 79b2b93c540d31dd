import torch, time
for mb in (64, 256, 1024, 2048):
    n = mb * 1024 * 1024 // 8
    h = torch.empty(n, dtype=torch.float64).pin_memory(); d = torch.empty(n, dtype=torch.float64, device="cuda")
    for name, src, dst in (("D2H", d, h), ("H2D", h, d)):
        dst.copy_(src, non_blocking=True); torch.cuda.synchronize()
        t0 = time.perf_counter()
        for _ in range(3): dst.copy_(src, non_blocking=True)
        torch.cuda.synchronize(); dt = (time.perf_counter() - t0) / 3
        print(f"{name} {mb:5d} MiB pinned: {n*8/dt*1e-9:6.1f} GB/s")
import subprocess
print(subprocess.run("nvidia-smi --query-gpu=pcie.link.gen.current,pcie.link.width.current,pcie.link.gen.max --format=csv; nproc; numactl -H 2>/dev/null | head -5", shell=True, capture_output=True, text=True).stdout)
