#!/usr/bin/env python3
"""Host <-> device copy bandwidth with pinned buffers, one GPU or all ranks of a torchrun launch at once.

  python tools/pcie_probe.py
  python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 tools/pcie_probe.py

Under torchrun every rank copies at the same time (barrier before, max over ranks of the device time), which gives the
aggregate the box's PCIe fabric and host memory sustain: the ceiling of the end-to-end arm of bench.py at N GPUs, whose
every step brings 1.92 GB of PCM per GPU back to the host. A second round pins the process to the CPUs next to its GPU
(/sys/bus/pci/devices/<bus id>/local_cpulist) before it allocates the pinned buffer, to show what NUMA placement is worth.
"""
import json
import os

import torch
import torch.distributed as dist

N = 1 << 30
REPS = 4


def local_cpus(dev):
    try:
        bus = torch.cuda.get_device_properties(dev)
        bus_id = f"{bus.pci_domain_id:04x}:{bus.pci_bus_id:02x}:{bus.pci_device_id:02x}.0"
        txt = open(f"/sys/bus/pci/devices/{bus_id}/local_cpulist").read().strip()
        cpus = set()
        for part in txt.split(","):
            a, _, b = part.partition("-")
            cpus.update(range(int(a), int(b or a) + 1))
        return bus_id, sorted(cpus & os.sched_getaffinity(0))
    except Exception as e:  # noqa: BLE001
        return repr(e), []


def measure(world, dev):
    h = torch.empty(N, dtype=torch.uint8).pin_memory()
    d = torch.empty(N, dtype=torch.uint8, device=dev)
    out = {}
    for name, fn in (("d2h", lambda: h.copy_(d, non_blocking=True)), ("h2d", lambda: d.copy_(h, non_blocking=True))):
        fn()
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(REPS):
            fn()
        e1.record()
        torch.cuda.synchronize()
        t = torch.tensor([e0.elapsed_time(e1)], device=dev)
        mine = REPS * N / (t.item() * 1e-3) / 1e9
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        out[name] = {"rank_gbs": mine, "aggregate_gbs": world * REPS * N / (t.item() * 1e-3) / 1e9}
    del h, d
    return out


def main():
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=dev)
    res = {"n_gpus": world, "host_cpus": len(os.sched_getaffinity(0)), "default": measure(world, dev)}
    bus_id, cpus = local_cpus(dev)
    res["bus_id"], res["local_cpus"] = bus_id, cpus
    if cpus:
        os.sched_setaffinity(0, cpus)
        res["pinned_to_local_cpus"] = measure(world, dev)
    if world > 1:
        allres = [None] * world
        dist.all_gather_object(allres, res)
        if rank == 0:
            os.write(2, (json.dumps({"per_rank": [{k: r[k] for k in ("bus_id", "local_cpus")} | {"d2h": r["default"]["d2h"]["rank_gbs"]} for r in allres]}) + "\n").encode())
        dist.destroy_process_group()
    if rank == 0:
        print(json.dumps(res), flush=True)


if __name__ == "__main__":
    main()
