"""Which host resource bounds the host-to-host (`e2e`) path when every GPU of the box copies at once?

    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 tools/pcie_scaling.py

Every rank copies 49 MB pinned host -> device and 52 MB device -> pinned host (the bench's e2e step), (a) alone, one
rank after the other, (b) all ranks at once, each direction alone and both together; then (c) the same with the
process pinned to the CPUs of the GPU's NUMA node before the pinned buffers are allocated (first touch).  Rank 0
prints one table: per-rank GB/s alone, per-rank and aggregate GB/s together, with and without NUMA placement.
"""
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import torch.distributed as dist

H2D, D2H = 49_152_000, 52_224_000


def numa_cpus_of_gpu(dev_index):
    try:
        props = torch.cuda.get_device_properties(dev_index)
        bus = "%04x:%02x:%02x.0" % (getattr(props, "pci_domain_id", 0), props.pci_bus_id, props.pci_device_id)
        node = int(open(f"/sys/bus/pci/devices/{bus}/numa_node").read())
        if node < 0:
            return None, node
        cpus = open(f"/sys/devices/system/node/node{node}/cpulist").read().strip()
        out = []
        for part in cpus.split(","):
            a, _, b = part.partition("-")
            out.extend(range(int(a), int(b or a) + 1))
        return out, node
    except Exception:
        return None, None


def measure(dev, reps=20):
    hin = torch.empty(H2D, dtype=torch.uint8).pin_memory()
    hout = torch.empty(D2H, dtype=torch.uint8).pin_memory()
    hin.fill_(1)
    din = torch.empty(H2D, dtype=torch.uint8, device=dev)
    dout = torch.empty(D2H, dtype=torch.uint8, device=dev)
    s1, s2 = torch.cuda.Stream(dev), torch.cuda.Stream(dev)

    def run(mode):
        torch.cuda.synchronize(dev)
        t0 = time.perf_counter()
        for _ in range(reps):
            if mode in ("h2d", "both"):
                with torch.cuda.stream(s1):
                    din.copy_(hin, non_blocking=True)
            if mode in ("d2h", "both"):
                with torch.cuda.stream(s2):
                    hout.copy_(dout, non_blocking=True)
        torch.cuda.synchronize(dev)
        dt = (time.perf_counter() - t0) / reps
        return ((H2D if mode != "d2h" else 0) + (D2H if mode != "h2d" else 0)) / dt / 1e9
    run("both")
    return run


def main():
    rank, world = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1))
    local = int(os.environ.get("LOCAL_RANK", 0))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("gloo")
    rows = {}
    for placement in ("default", "numa-local"):
        cpus, node = numa_cpus_of_gpu(local)
        if placement == "numa-local":
            if not cpus:
                rows[placement] = None
                continue
            os.sched_setaffinity(0, cpus)
        run = measure(dev)
        res = {}
        for mode in ("h2d", "d2h", "both"):
            alone = 0.0
            for r in range(world):                      # one rank after the other
                if world > 1:
                    dist.barrier()
                if r == rank:
                    alone = run(mode)
            if world > 1:
                dist.barrier()
            together = run(mode)                        # all ranks at once
            res[mode] = (alone, together)
        rows[placement] = (res, node)
    gathered = [None] * world
    if world > 1:
        dist.all_gather_object(gathered, rows)
    else:
        gathered = [rows]
    if rank == 0:
        for placement in ("default", "numa-local"):
            print(f"== host buffers: {placement}")
            if gathered[0][placement] is None:
                print("   (no NUMA information for the GPUs on this box)")
                continue
            for mode in ("h2d", "d2h", "both"):
                alone = [g[placement][0][mode][0] for g in gathered]
                tog = [g[placement][0][mode][1] for g in gathered]
                print(f"   {mode:5s} alone per rank: " + " ".join(f"{v:5.1f}" for v in alone) +
                      f" | together per rank: " + " ".join(f"{v:5.1f}" for v in tog) + f" | aggregate {sum(tog):6.1f} GB/s")
            print("   NUMA nodes of the GPUs:", [g[placement][1] for g in gathered])
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
