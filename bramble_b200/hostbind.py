"""Host-side placement for one-process-per-GPU runs: keep a rank's threads (and the pinned staging pages they first
touch) on the NUMA node its GPU hangs off.  Plain sysfs + sched_setaffinity; no libnuma in this image."""
import os


def _parse_cpulist(s):
    cpus = set()
    for part in s.strip().split(","):
        if not part:
            continue
        a, _, b = part.partition("-")
        cpus.update(range(int(a), int(b or a) + 1))
    return cpus


def gpu_numa_node(device):
    """NUMA node of CUDA device `device` (from its PCI address), or None when the platform does not say."""
    try:
        import torch
        p = torch.cuda.get_device_properties(device)
        addr = "%04x:%02x:%02x.0" % (p.pci_domain_id, p.pci_bus_id, p.pci_device_id)
        with open(f"/sys/bus/pci/devices/{addr}/numa_node") as fh:
            node = int(fh.read().strip())
        return node if node >= 0 else None
    except Exception:
        return None


def describe(device):
    node = gpu_numa_node(device)
    return {"numa_node": node, "cpus_allowed": len(os.sched_getaffinity(0)), "bound": False}


def bind_to_gpu_node(device, local_rank=None, local_world=None):
    """Restrict this process to the CPUs of the GPU's NUMA node (intersected with what it may already use).  On a
    platform without NUMA information the process is left where it is -- unless local_rank / local_world are
    given: the allowed CPUs are then dealt out evenly, rank r of n taking the r-th slice, so that the ranks' host
    threads (each rank's main thread sits in the launch -> synchronise loop of every step) do not preempt one
    another.  Returns what was done."""
    info = describe(device)
    node = info["numa_node"]
    if node is None:
        if local_rank is not None and local_world and local_world > 1:
            cpus = sorted(os.sched_getaffinity(0))
            per = len(cpus) // local_world
            if per >= 2:
                mine = set(cpus[local_rank * per:(local_rank + 1) * per])
                try:
                    os.sched_setaffinity(0, mine)
                    info.update(cpus_allowed=len(mine), bound=True, how="even slice of the allowed CPUs")
                except OSError:
                    pass
        return info
    try:
        with open(f"/sys/devices/system/node/node{node}/cpulist") as fh:
            cpus = _parse_cpulist(fh.read()) & os.sched_getaffinity(0)
        if cpus:
            os.sched_setaffinity(0, cpus)
            info.update(cpus_allowed=len(cpus), bound=True)
    except Exception:
        pass
    return info
