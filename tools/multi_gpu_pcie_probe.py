"""How much pinned-host <-> device bandwidth do N GPUs of one box get AT THE SAME TIME, and does it matter which
NUMA node the pinned buffers and the issuing thread live on?

    python tools/multi_gpu_pcie_probe.py 0,1,2,3 [bind]

One child process per GPU; each moves 12 MiB H2D + 6 MiB D2H per step on two streams (the shape of one bench step)
for about two seconds, all children starting on a shared wall-clock deadline.  With `bind` a child first restricts
itself to the CPUs of its GPU's NUMA node (sysfs), so its pinned buffers are first-touched there.
"""
import os
import subprocess
import sys
import time


def gpu_numa_cpus(index):
    """CPUs of the NUMA node the GPU hangs off, or None when the box does not say."""
    import torch
    p = torch.cuda.get_device_properties(index)
    bdf = "%04x:%02x:%02x.0" % (p.pci_domain_id, p.pci_bus_id, p.pci_device_id)
    try:
        node = int(open(f"/sys/bus/pci/devices/{bdf}/numa_node").read())
        if node < 0:
            return bdf, node, None
        cpus = set()
        for part in open(f"/sys/devices/system/node/node{node}/cpulist").read().strip().split(","):
            a, _, b = part.partition("-")
            cpus.update(range(int(a), int(b or a) + 1))
        return bdf, node, cpus
    except OSError:
        return bdf, None, None


def child(index, bind, deadline):
    import torch
    torch.cuda.set_device(index)
    bdf, node, cpus = gpu_numa_cpus(index)
    note = f"gpu {index} {bdf} node {node}"
    if bind and cpus:
        allowed = cpus & os.sched_getaffinity(0)
        if allowed:
            os.sched_setaffinity(0, allowed)
            note += f" bound to {len(allowed)} cpus"
        else:
            note += " (node cpus not in this process's affinity mask)"
    n = 12 << 20
    h = torch.empty(n, dtype=torch.uint8).pin_memory()
    h.fill_(1)
    d = torch.empty(n, dtype=torch.uint8, device="cuda")
    h2 = torch.empty(n // 2, dtype=torch.uint8).pin_memory()
    h2.fill_(1)
    d2 = torch.empty(n // 2, dtype=torch.uint8, device="cuda")
    s1, s2 = torch.cuda.Stream(), torch.cuda.Stream()
    for _ in range(5):
        d.copy_(h, non_blocking=True)
    torch.cuda.synchronize()
    while time.time() < deadline:
        time.sleep(0.001)
    reps = 0
    t = time.perf_counter()
    while time.perf_counter() - t < 2.0:
        for _ in range(20):
            with torch.cuda.stream(s1):
                d.copy_(h, non_blocking=True)
            with torch.cuda.stream(s2):
                h2.copy_(d2, non_blocking=True)
        torch.cuda.synchronize()
        reps += 20
    dt = time.perf_counter() - t
    print(f"{note}: H2D {n * reps / dt / 1e9:.1f} GB/s + D2H {n / 2 * reps / dt / 1e9:.1f} GB/s", flush=True)


if __name__ == "__main__":
    if len(sys.argv) >= 2 and sys.argv[1] == "--child":
        child(int(sys.argv[2]), sys.argv[3] == "1", float(sys.argv[4]))
    else:
        gpus = [int(x) for x in sys.argv[1].split(",")]
        bind = len(sys.argv) > 2 and sys.argv[2] == "bind"
        deadline = time.time() + 15.0
        ps = [subprocess.Popen([sys.executable, __file__, "--child", str(g), "1" if bind else "0", str(deadline)])
              for g in gpus]
        for p in ps:
            p.wait()
