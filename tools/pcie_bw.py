import torch, time
x = torch.empty(64 << 20, dtype=torch.uint8).pin_memory()
d = torch.empty(64 << 20, dtype=torch.uint8, device="cuda")
for _ in range(3): d.copy_(x, non_blocking=True)
torch.cuda.synchronize()
t0 = time.perf_counter()
for _ in range(20): d.copy_(x, non_blocking=True)
torch.cuda.synchronize()
dt = (time.perf_counter() - t0) / 20
print("H2D pinned 64 MiB: %.3f ms, %.1f GB/s" % (dt * 1e3, 64 * 1.048576e-3 / dt / 1e0 * 1e0))
y = torch.empty(8 << 20, dtype=torch.uint8).pin_memory(); e = torch.empty(8 << 20, dtype=torch.uint8, device="cuda")
torch.cuda.synchronize(); t0 = time.perf_counter()
for _ in range(20): y.copy_(e, non_blocking=True)
torch.cuda.synchronize(); dt = (time.perf_counter() - t0) / 20
print("D2H pinned 8 MiB: %.3f ms" % (dt * 1e3))
import subprocess
print(subprocess.run(["nvidia-smi", "--query-gpu=pcie.link.gen.current,pcie.link.width.current", "--format=csv,noheader"], capture_output=True, text=True).stdout)
