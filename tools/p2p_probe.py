"""Topology / peer-access probe for the 2-GPU box: is cuda:0 <-> cuda:1 NVLink P2P, and how fast?"""
import subprocess, torch, time
print(subprocess.run(["nvidia-smi", "topo", "-m"], capture_output=True, text=True).stdout)
print("can_access_peer 0->1:", torch.cuda.can_device_access_peer(0, 1))
a = torch.empty(256 << 20, dtype=torch.uint8, device="cuda:0")
b = torch.empty(256 << 20, dtype=torch.uint8, device="cuda:1")
for _ in range(3): b.copy_(a)
torch.cuda.synchronize(0); torch.cuda.synchronize(1)
t0 = time.perf_counter()
for _ in range(10): b.copy_(a)
torch.cuda.synchronize(0); torch.cuda.synchronize(1)
print("D2D 0->1: %.1f GB/s" % (10 * a.numel() / (time.perf_counter() - t0) / 1e9))
s = a[:4096]; d = b[:4096]
torch.cuda.synchronize(0); torch.cuda.synchronize(1)
t0 = time.perf_counter()
for _ in range(200): d.copy_(s)
torch.cuda.synchronize(0); torch.cuda.synchronize(1)
print("4 KiB copy latency: %.1f us" % ((time.perf_counter() - t0) / 200 * 1e6))
