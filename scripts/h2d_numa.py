"""Concurrent H2D bandwidth of all GPUs of the box, one process per GPU (torchrun), with and without binding each
rank to the NUMA node of its GPU before the pinned buffer is allocated."""
import os, sys, time, glob
import torch
rank = int(os.environ.get("LOCAL_RANK", "0"))
bind = len(sys.argv) > 1 and sys.argv[1] == "bind"
torch.cuda.set_device(rank)
prop_bus = torch.cuda.get_device_properties(rank).pci_bus_id if hasattr(torch.cuda.get_device_properties(rank), "pci_bus_id") else None
import subprocess
bus = subprocess.run(["nvidia-smi", "--query-gpu=pci.bus_id", "--format=csv,noheader", "-i", str(rank)], capture_output=True, text=True).stdout.strip().lower()
node = -1
try:
    node = int(open(f"/sys/bus/pci/devices/{bus[4:] if bus.startswith('0000') and len(bus) > 12 else bus}/numa_node").read())
except Exception as e:
    node = -2
cpus = None
if bind and node >= 0:
    txt = open(f"/sys/devices/system/node/node{node}/cpulist").read().strip()
    cpus = set()
    for part in txt.split(","):
        a, _, b = part.partition("-")
        cpus.update(range(int(a), int(b or a) + 1))
    os.sched_setaffinity(0, cpus)
n = 1 << 30
host = torch.empty(n, dtype=torch.uint8, pin_memory=True)
host.fill_(1)                      # first touch on the bound node
dev = torch.empty(n, dtype=torch.uint8, device="cuda")
import torch.distributed as dist
dist.init_process_group("nccl", device_id=torch.device(f"cuda:{rank}"))
for _ in range(2):
    dev.copy_(host, non_blocking=True)
torch.cuda.synchronize(); dist.barrier()
t0 = time.perf_counter()
for _ in range(5):
    dev.copy_(host, non_blocking=True)
torch.cuda.synchronize()
dt = time.perf_counter() - t0
print(f"rank {rank} bus {bus} numa {node} bind {bind} cpus {len(cpus) if cpus else len(os.sched_getaffinity(0))} H2D {5 * n / dt / 1e9:.1f} GB/s", file=sys.stderr, flush=True)
dist.destroy_process_group()
