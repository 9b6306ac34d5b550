"""Host -> device copy rate with every rank of the box copying at once (run under torchrun): what bounds bench.py's `e2e` at N > 1.
Each rank copies a pinned 2 GiB buffer to its GPU REPS times after a barrier; rank 0 prints the per-rank rates and the aggregate
(all bytes / slowest rank's time).  usage: torchrun --nproc-per-node N scripts/pcie_concurrent.py"""
import os

import torch
import torch.distributed as dist

rank, world, local = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1)), int(os.environ.get("LOCAL_RANK", 0))
torch.cuda.set_device(local)


def numa_cpus_of_gpu(index: int):
    """CPUs of the NUMA node the GPU's PCIe slot hangs off (None when the platform does not say)."""
    try:
        bus = torch.cuda.get_device_properties(index).pci_bus_id
        dom = torch.cuda.get_device_properties(index).pci_domain_id
        dev = torch.cuda.get_device_properties(index).pci_device_id
        path = f"/sys/bus/pci/devices/{dom:04x}:{bus:02x}:{dev:02x}.0/numa_node"
        node = int(open(path).read())
        if node < 0:
            return None, node
        cpus = open(f"/sys/devices/system/node/node{node}/cpulist").read().strip()
        out = set()
        for part in cpus.split(","):
            a, _, b = part.partition("-")
            out.update(range(int(a), int(b or a) + 1))
        return out, node
    except Exception as e:  # noqa: BLE001
        return None, repr(e)


cpus, node = numa_cpus_of_gpu(local)
if os.environ.get("BIND") and cpus:
    os.sched_setaffinity(0, cpus)  # before the pinned buffer is first touched: its pages land on this node
if world > 1:
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
BYTES, REPS = 2 << 30, 5
h = torch.empty(BYTES, dtype=torch.uint8).pin_memory()
h.fill_(1)
d = torch.empty(BYTES, dtype=torch.uint8, device="cuda")
d.copy_(h, non_blocking=True)
torch.cuda.synchronize()
if world > 1:
    dist.barrier()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(REPS):
    d.copy_(h, non_blocking=True)
e1.record()
torch.cuda.synchronize()
ms = e0.elapsed_time(e1)
rate = BYTES * REPS / ms / 1e6
rates = [None] * world
if world > 1:
    dist.all_gather_object(rates, (rate, ms))
else:
    rates = [(rate, ms)]
if rank == 0:
    slowest = max(m for _, m in rates)
    print(f"bind={bool(os.environ.get('BIND'))} gpu0 numa node {node} cpus {len(cpus) if cpus else None}")
    print(f"H2D pinned, {world} ranks at once: per rank {[round(r, 1) for r, _ in rates]} GB/s; aggregate {world * BYTES * REPS / slowest / 1e6:.1f} GB/s "
          f"({os.cpu_count()} host cores)")
if world > 1:
    dist.destroy_process_group()
