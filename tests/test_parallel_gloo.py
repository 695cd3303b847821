"""CPU test of the N>1 path: two `gloo` processes each own a contiguous slab of ensemble members (pam_b200.parallel
.shard_range), step their slab independently (the oracle stands in for the GPU kernels on this CPU-only host: the
sharding and the collective are what is under test) and combine the only collective of the path, the global mass
diagnostic. The sharded result must equal the single-process result member for member."""
import os
import socket
import sys

import numpy as np
import pytest
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
NENS = 5           # odd on purpose: slabs of 3 and 2 members


def _free_port():
    s = socket.socket(); s.bind(("127.0.0.1", 0)); p = s.getsockname()[1]; s.close()
    return p


def _worker(rank, world, port, out):
    sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
    os.environ["MASTER_ADDR"] = "127.0.0.1"; os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    import helpers as H
    from pam_b200 import synthetic as syn
    from pam_b200.parallel import shard_range, global_mass, global_min_dt
    e0, e1 = shard_range(NENS, rank, world)
    case = syn.make_case(12, 1, 10, e1 - e0, grid="tanh", ens_offset=e0)
    o = H.make_oracle(case, 5, crm_dt=2.0)
    o.declare_hydrostatic()
    dt_local = o.compute_time_step()
    dt_glob = global_min_dt(dt_local)
    o.time_step()
    gm = global_mass(o.compute_mass(), NENS)
    np.savez(os.path.join(out, "rank%d.npz" % rank), e0=e0, e1=e1, dt_local=dt_local, dt_glob=dt_glob, gm=gm,
             temp=o.get_field("temp"), uvel=o.get_field("uvel"))
    dist.barrier()
    dist.destroy_process_group()


def test_two_rank_ensemble_sharding_matches_single_process(tmp_path):
    world = 2
    mp.spawn(_worker, args=(world, _free_port(), str(tmp_path)), nprocs=world, join=True)
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import helpers as H
    from pam_b200 import synthetic as syn
    case = syn.make_case(12, 1, 10, NENS, grid="tanh")
    o = H.make_oracle(case, 5, crm_dt=2.0)
    o.declare_hydrostatic()
    dt_all = o.compute_time_step()
    o.time_step()
    mass = o.compute_mass()
    ranks = [np.load(os.path.join(str(tmp_path), "rank%d.npz" % r)) for r in range(world)]
    assert int(ranks[0]["e0"]) == 0 and int(ranks[0]["e1"]) == int(ranks[1]["e0"]) and int(ranks[1]["e1"]) == NENS
    for r in ranks:
        e0, e1 = int(r["e0"]), int(r["e1"])
        # CRMs are independent: a member's result does not depend on which rank owns it
        assert np.array_equal(r["temp"], o.get_field("temp")[..., e0:e1])
        assert np.array_equal(r["uvel"], o.get_field("uvel")[..., e0:e1])
        # the all-reduced diagnostics agree on every rank and with the single-process values
        assert np.allclose(r["gm"], mass.mean(axis=1), rtol=1e-14, atol=0)
        assert float(r["dt_glob"]) == min(float(x["dt_local"]) for x in ranks)
    assert abs(min(float(x["dt_local"]) for x in ranks) - dt_all) <= 1e-15 * dt_all


def test_bind_to_gpu_numa_node_on_a_fake_topology(tmp_path, monkeypatch):
    """Rank placement for the host-buffer path: the process is pinned to the CPUs of its GPU's NUMA node (sysfs walk),
    and left alone when the topology is not visible."""
    import types
    import torch
    from pam_b200 import parallel as P
    props = types.SimpleNamespace(pci_domain_id=0, pci_bus_id=0x1b, pci_device_id=0)
    monkeypatch.setattr(torch.cuda, "get_device_properties", lambda i: props)
    allowed = sorted(os.sched_getaffinity(0))
    dev = tmp_path / "bus/pci/devices/0000:1b:00.0"
    dev.mkdir(parents=True)
    (dev / "numa_node").write_text("1\n")
    node = tmp_path / "devices/system/node/node1"
    node.mkdir(parents=True)
    (node / "cpulist").write_text("%d\n" % allowed[0])
    try:
        assert P.bind_to_gpu_numa_node(0, sysfs=str(tmp_path)) == {"node": 1, "cpus": 1}
        assert os.sched_getaffinity(0) == {allowed[0]}
    finally:
        os.sched_setaffinity(0, allowed)
    (dev / "numa_node").write_text("-1\n")
    assert P.bind_to_gpu_numa_node(0, sysfs=str(tmp_path)) is None
    (dev / "numa_node").write_text("1\n")
    (node / "cpulist").write_text("100000-100003\n")                     # CPUs outside this process's mask
    assert P.bind_to_gpu_numa_node(0, sysfs=str(tmp_path)) is None
    assert P.bind_to_gpu_numa_node(0, sysfs=str(tmp_path / "nowhere")) is None
    assert os.sched_getaffinity(0) == set(allowed)
    assert P._parse_cpulist("0-3,8,10-11\n") == {0, 1, 2, 3, 8, 10, 11}
