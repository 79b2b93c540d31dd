import torch, time
torch.backends.cuda.matmul.allow_tf32=False
for n in (2048,4096,8192):
    a=torch.randn(n,n,dtype=torch.float64,device='cuda'); b=torch.randn(n,n,dtype=torch.float64,device='cuda')
    for _ in range(2): c=a@b
    torch.cuda.synchronize()
    best=1e9
    for _ in range(5):
        e0=torch.cuda.Event(enable_timing=True); e1=torch.cuda.Event(enable_timing=True)
        e0.record(); c=a@b; e1.record(); torch.cuda.synchronize()
        best=min(best,e0.elapsed_time(e1))
    print(f"torch fp64 matmul n={n}: {2*n**3/best*1e-9:.2f} TFLOP/s ({best:.3f} ms)")
# batched 512
a=torch.randn(148,512,512,dtype=torch.float64,device='cuda'); b=torch.randn(148,512,1152,dtype=torch.float64,device='cuda')
for _ in range(2): c=a@b
torch.cuda.synchronize()
e0=torch.cuda.Event(enable_timing=True); e1=torch.cuda.Event(enable_timing=True)
e0.record(); c=a@b; e1.record(); torch.cuda.synchronize()
ms=e0.elapsed_time(e1); print(f"bmm 148x512x512x1152: {2*148*512*512*1152/ms*1e-9:.2f} TFLOP/s ({ms:.3f} ms)")
import subprocess
print(subprocess.run("nvidia-smi --query-gpu=clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active --format=csv",shell=True,capture_output=True,text=True).stdout)
