"""Ensemble sharding across the GPUs of one node and the global diagnostics that go with it.

The AWFL step has no data-path exchange between ranks: every stencil runs along k, j or i at a FIXED ensemble member
(SURVEY.md 8e), so each rank owns a contiguous slab of members with its own coupler + handle -- the way E3SM-MMF places
a fixed number of CRMs per MPI rank. The only collective is the global mass/energy diagnostic, the multi-GPU form of
Dycore::compute_mass (dynamics/awfl/Dycore.h:37-60): one all-reduce of (num_tracers+2) doubles (NCCL on GPUs, gloo in
the CPU tests).
"""
import numpy as np
import torch
import torch.distributed as dist


def shard_range(nens_global, rank, world):
    """Contiguous member slab [e0, e1) of `rank`; slabs differ by at most one member and cover [0, nens_global)."""
    if not (0 <= rank < world):
        raise ValueError("rank %d outside world of %d" % (rank, world))
    base, rem = divmod(int(nens_global), int(world))
    e0 = rank * base + min(rank, rem)
    return e0, e0 + base + (1 if rank < rem else 0)


def global_mass(local_mass, nens_global, group=None, device=None):
    """local_mass: (nvars, nens_local) per-member means of q*dz (Dycore::compute_mass). Returns the (nvars,) ensemble
    means over ALL ranks' members: sum over local members, all-reduce(SUM), divide by the global member count."""
    lm = np.asarray(local_mass, dtype=np.float64)
    s = torch.from_numpy(lm.sum(axis=1).copy())
    if dist.is_available() and dist.is_initialized() and dist.get_world_size(group) > 1:
        if dist.get_backend(group) == "nccl":
            s = s.to(device if device is not None else torch.device("cuda", torch.cuda.current_device()))
        dist.all_reduce(s, op=dist.ReduceOp.SUM, group=group)
    return (s.cpu().numpy()) / float(nens_global)


def global_min_dt(local_dt, group=None, device=None):
    """Optional uniform sub-cycle count across ranks: min of the per-rank CFL time steps (each rank of the reference
    computes its own, Dycore.h:141; benchmarks may want one)."""
    t = torch.tensor([float(local_dt)], dtype=torch.float64)
    if dist.is_available() and dist.is_initialized() and dist.get_world_size(group) > 1:
        if dist.get_backend(group) == "nccl":
            t = t.to(device if device is not None else torch.device("cuda", torch.cuda.current_device()))
        dist.all_reduce(t, op=dist.ReduceOp.MIN, group=group)
    return float(t.item())


def _parse_cpulist(text):
    cpus = set()
    for part in text.strip().split(","):
        if not part:
            continue
        a, _, b = part.partition("-")
        cpus.update(range(int(a), int(b or a) + 1))
    return cpus


def bind_to_gpu_numa_node(device_index, sysfs="/sys"):
    """Pin the calling process to the CPUs of the NUMA node its GPU hangs off, BEFORE it allocates host buffers: the
    first touch then places the (pinned) coupler fields in that node's memory, so the H2D/D2H streams of the host-buffer
    entry point do not cross the socket interconnect -- with 8 ranks x 2 x 55 GB/s that is what decides the e2e rate.
    Returns {"node": n, "cpus": count} or None when the topology is not visible (no sysfs entry, single node, or the
    node's CPUs are outside this process's affinity mask). E3SM places MPI ranks the same way through its launcher."""
    import os
    try:
        p = torch.cuda.get_device_properties(device_index)
        bdf = "%04x:%02x:%02x.0" % (p.pci_domain_id, p.pci_bus_id, p.pci_device_id)
        with open(os.path.join(sysfs, "bus/pci/devices", bdf, "numa_node")) as fh:
            node = int(fh.read().strip())
        if node < 0:
            return None
        with open(os.path.join(sysfs, "devices/system/node/node%d/cpulist" % node)) as fh:
            cpus = _parse_cpulist(fh.read())
        cpus &= os.sched_getaffinity(0)
        if not cpus:
            return None
        os.sched_setaffinity(0, cpus)
        return {"node": node, "cpus": len(cpus)}
    except (OSError, ValueError, AttributeError):
        return None
