#!/usr/bin/env python3
"""bench.py -- AWFL dycore time-step throughput (cell-updates/s) on N B200s of one node.

  python bench.py --gpus N --steps K --warmup W              # the CUDA path (this repo)
  python bench.py --impl reference --gpus N --steps K ...    # the reference's own CPU code (oracle/_ref, OpenMP)

A "step" is one Dycore::timeStep over the whole ensemble batch (= `ncycles` SSPRK3 sub-cycles); one cell-update is one
(k,j,i,iens) cell advanced by one sub-cycle (3 Runge-Kutta stages), SURVEY.md 8(d). Workload at N=1 is BASELINE.json
configs[1]: MMF-style batched 2-D CRMs 64x58, nens=4096, WENO ord=5, one tracer (water vapour), supercell sounding with
per-member wind offsets and a seeded temperature perturbation (pam_b200/synthetic.py). N>1: every rank owns its own
4096 members (weak scaling, the way MMF places a fixed number of CRMs per rank); no data-path collective exists
(CRMs are independent) -- NCCL only carries the timing max and the global mass diagnostic.

The same JSON line carries, under "strong_c3", BASELINE.json configs[2] -- 3-D CRMs 32x32x58 with nens_global = 1024 FIXED,
ensemble-sharded over the N ranks (1024 / 512 / 256 / 128 members per GPU): the strong-scaling curve the >= 7x target is
read against. `value` (the driver's headline and efficiency input) stays the weak C2 line.

Prints ONE JSON line (rank 0).
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

WORKLOADS = {
    # name: (nz, ny, nx, nens_per_gpu, ord, tracer set)
    "C2": dict(nz=58, ny=1, nx=64, nens=4096, ord=5, nt=1, desc="2-D 64x58, nens=4096/GPU, ord=5, NT=1"),
    "C3": dict(nz=58, ny=32, nx=32, nens=128, ord=5, nt=1, desc="3-D 32x32x58, nens=128/GPU, ord=5, NT=1"),
    "C4": dict(nz=58, ny=1, nx=128, nens=2048, ord=5, nt=1, desc="2-D 128x58, nens=2048/GPU, ord=5, NT=1"),
    "C5": dict(nz=58, ny=1, nx=64, nens=4096, ord=5, nt=10, desc="2-D 64x58, nens=4096/GPU, ord=5, NT=10 (P3+SHOC tracers)"),
    # BASELINE.json configs[0]: the reference's own CPU-runnable case -- the standalone driver's supercell sounding (Dycore.h:1096-1276),
    # nx=64 nz=50 nens=1, the dycore's default constants (Dycore.h:871-876)
    "C1": dict(nz=50, ny=1, nx=64, nens=1, ord=5, nt=1, init="supercell", default_consts=True,
               desc="2-D 64x50, nens=1, ord=5, NT=1, supercell sounding (standalone-driver case)"),
    "tiny": dict(nz=16, ny=1, nx=16, nens=64, ord=5, nt=1, desc="2-D 16x16, nens=64 (plumbing check)"),
}
CRM_DT = 1.0


def build_case(w, ens_offset=0, nens=None):
    """Synthetic fields are generated for a 32-member tile and tiled along the ensemble axis with the per-member
    wind offset re-applied, so that set-up of 15M-cell fields stays cheap; values still differ member to member."""
    from pam_b200 import synthetic as syn
    nens = nens or w["nens"]
    tr = syn.TRACERS_NT1 if w["nt"] == 1 else syn.TRACERS_P3_SHOC
    base_n = min(nens, 68)   # 68 = 4*17: covers every residue of the %17, %7, %5, %3 member variations
    case = syn.make_case(w["nz"], w["ny"], w["nx"], base_n, tracers=tr, grid="equal", ens_offset=ens_offset)
    if base_n < nens:
        reps = (nens + base_n - 1) // base_n
        for k, v in case["fields"].items():
            case["fields"][k] = np.ascontiguousarray(np.tile(v, (1, 1, 1, reps))[..., :nens])
        case["nens"] = nens
    if w.get("default_consts"):
        case["consts"] = dict(R_d=287., cp_d=1003., R_v=461., cp_v=1859., p0=1.e5, grav=9.81)     # Dycore.h:871-876
    return case


class ClockSampler:
    """nvidia-smi clocks / throttle reasons during the timed region (B200_PROFILING.md)."""

    def __init__(self, index):
        self.index, self.rows, self.proc = index, [], None

    def start(self):
        q = "clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown," \
            "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap"
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), "--query-gpu=" + q, "--format=csv,noheader,nounits",
                                          "-lms", "100"], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.th = threading.Thread(target=self._read, daemon=True)
            self.th.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([x.strip() for x in line.split(",")])

    def stop(self):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except Exception:
            pass
        sm = [float(r[0]) for r in self.rows if len(r) >= 7 and r[0].replace(".", "").isdigit()]
        mx = [float(r[1]) for r in self.rows if len(r) >= 7 and r[1].replace(".", "").isdigit()]
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = [n for i, n in enumerate(names) if any(len(r) >= 7 and r[3 + i].lower().startswith("active") for r in self.rows)]
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None, "reasons": reasons,
                "samples": len(sm)}


def cpu_reference_rate(w, seconds_target=12.0, warm=1, steps=None, nens=None):
    """Times the reference's own CPU implementation (oracle/_ref, OpenMP build of the unmodified reference) on a bounded
    sample of the workload: same grid, order, tracers and inputs, fewer ensemble members. Returns the cpu_baseline dict
    and the per-step times."""
    from oracle import reflib
    cores = int(os.environ.get("AWFL_REF_THREADS", "0")) or len(os.sched_getaffinity(0)) or os.cpu_count() or 1
    # all the host threads the box offers: torchrun exports OMP_NUM_THREADS=1 to its workers, which is not what the
    # reference arm is asked to measure, so the value is overridden here (before libgomp reads it) and again through
    # omp_set_num_threads in case an OpenMP runtime is already initialised in this process
    os.environ["OMP_NUM_THREADS"] = str(cores)
    os.environ.setdefault("OMP_PROC_BIND", "close")
    try:
        import ctypes
        ctypes.CDLL("libgomp.so.1").omp_set_num_threads(cores)
    except OSError:
        pass
    kind = "reference"
    if not reflib.available(w["ord"], omp=True):
        raise RuntimeError("oracle/_ref/libpam_ref_ord%d_omp.so is missing (build it with __graft_entry__.build() where "
                           "/root/reference exists)" % w["ord"])
    # bounded sample: 8 members per host thread (a workload with fewer members, like C1, runs exactly as specified)
    nens = nens or (w["nens"] if w["nens"] <= 8 else max(8, min(w["nens"], 8 * cores)))
    case = build_case(w, nens=nens)
    r = reflib.RefDycore(w["nz"], w["ny"], w["nx"], nens, case["xlen"], case["ylen"], case["zint"], tracers=case["tracers"],
                         consts=case["consts"], crm_dt=CRM_DT, ord=w["ord"], omp=True)
    if w.get("init") == "supercell":
        r.init_supercell()
    else:
        r.set_fields(case["fields"])
    r.declare_hydrostatic_raw(); r.declare_hydrostatic_raw()    # twice: SURVEY.md F2/F4
    dt = r.compute_time_step()
    ncycles = int(np.ceil(CRM_DT / dt))
    for _ in range(warm):
        r.time_steps_raw(1)
    t1 = r.time_steps_raw(1)
    if steps is None:
        steps = int(max(2, min(50, seconds_target / max(t1, 1e-3))))
    times = [r.time_steps_raw(1) for _ in range(steps)]
    upd = w["nz"] * w["ny"] * w["nx"] * nens * ncycles
    rate = upd * len(times) / sum(times)
    cb = {"value": rate, "unit": "cell-updates/s", "cores": cores, "kind": kind,
          "sample": "%s with nens=%d (of %d), ncycles=%d, %d timed Dycore::timeStep calls of the unmodified reference "
                    "(prebuilt in the build container with g++ -O3 -march=x86-64-v3 -- not -march=native, so that the library also runs on this "
                    "host -- YAKL OpenMP backend, OMP_NUM_THREADS=%d)" % (w["desc"], nens, w["nens"], ncycles, len(times), cores)}
    return cb, times, ncycles, upd


def make_config(args, w, world, ncycles):
    """`config` of the JSON line: names the WORKLOAD (the same for both arms; what a bounded sample ran is said in cpu_baseline.sample)."""
    ncell = w["nz"] * w["ny"] * w["nx"] * w["nens"]
    nf = 5 + (1 if w["nt"] == 1 else 10)
    return {"workload": args.workload + ": " + w["desc"], "nens_global": w["nens"] * world, "crm_dt": CRM_DT, "ncycles": ncycles,
            "cell_updates_per_step": float(ncell) * ncycles * world,
            "l2": "inputs larger than L2: %d fields x %.0f MB per GPU, no flush needed" % (nf, 8 * ncell / 1e6),
            "parallelism": "ensemble-sharded x%d, no data-path collective" % world,
            "lines": "`value` / `e2e`: weak scaling of this workload (fixed members per GPU); `strong_c3`: strong scaling of "
                     "BASELINE.json configs[2] (3-D 32x32x58, nens_global=1024 fixed)"}


def run_reference_arm(args, w):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    cb, times, ncycles, upd = cpu_reference_rate(w, warm=max(1, args.warmup), steps=args.steps)
    ms = 1e3 * sum(times) / len(times)
    line = {"impl": "reference", "metric": "AWFL dycore cell-updates/sec", "value": cb["value"], "unit": "cell-updates/s",
            "n_gpus": args.gpus, "steps": len(times), "warmup": max(1, args.warmup), "ms_per_step": ms, "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": make_config(args, w, max(1, args.gpus), ncycles),
            "cpu_baseline": cb,
            "e2e": {"value": cb["value"], "unit": "cell-updates/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    print(json.dumps(line))


def host_link_rate(dev, world, mbytes=512):
    """Pinned host <-> device copy rate of THIS rank with both directions busy (GB/s per direction), the way the host-buffer
    entry point uses the link; every rank measures at the same time, so at N > 1 the number includes the contention for the
    box's host fabric. Returns (this rank's rate, min over ranks, sum over ranks)."""
    import torch
    import torch.distributed as dist
    n = mbytes * (1 << 20) // 8
    h_in, h_out = torch.empty(n, dtype=torch.float64).pin_memory(), torch.empty(n, dtype=torch.float64).pin_memory()
    d_in, d_out = torch.empty(n, dtype=torch.float64, device=dev), torch.zeros(n, dtype=torch.float64, device=dev)
    s1, s2 = torch.cuda.Stream(dev), torch.cuda.Stream(dev)
    best = 0.0
    for it in range(3):
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize(dev)
        e0, e1, e2 = (torch.cuda.Event(enable_timing=True) for _ in range(3))
        e0.record()
        s1.wait_event(e0); s2.wait_event(e0)
        with torch.cuda.stream(s1):
            d_in.copy_(h_in, non_blocking=True); e1.record()
        with torch.cuda.stream(s2):
            h_out.copy_(d_out, non_blocking=True); e2.record()
        torch.cuda.synchronize(dev)
        ms = max(e0.elapsed_time(e1), e0.elapsed_time(e2))
        if it > 0:
            best = max(best, 8.0 * n / (ms * 1e-3) / 1e9)
    t = torch.tensor([best, best], dtype=torch.float64, device=dev)
    lo, tot = best, best
    if world > 1:
        tmin = t[:1].clone(); dist.all_reduce(tmin, op=dist.ReduceOp.MIN)
        tsum = t[1:].clone(); dist.all_reduce(tsum, op=dist.ReduceOp.SUM)
        lo, tot = float(tmin.item()), float(tsum.item())
    return best, lo, tot


def run_workload(w, args, rank, world, dev, nens_global, steps, warmup, do_e2e, do_profile, numa=None, link=None):
    """Set up this rank's shard of `nens_global` members of workload `w`, check a few members against the oracle, time
    `steps` device-resident steps (CUDA events, max over ranks) and the same step through host buffers."""
    import torch
    import torch.distributed as dist
    from pam_b200 import PamCoupler, Dycore
    from pam_b200.parallel import shard_range, global_mass

    e0, e1 = shard_range(nens_global, rank, world)
    nloc = e1 - e0
    case = build_case(w, ens_offset=e0, nens=nloc)
    c = PamCoupler(dev)
    c.set_option("crm_dt", CRM_DT)
    for k in ("R_d", "cp_d", "R_v", "cp_v", "p0", "grav"):
        c.set_option(k, case["consts"][k])
    c.allocate_coupler_state(w["nz"], w["ny"], w["nx"], nloc)
    c.set_grid(case["xlen"], case["ylen"], case["zint"])
    for name, pos, adds in case["tracers"]:
        c.add_tracer(name, name, pos, adds)
    d = Dycore(w["ord"])
    d.init(c)
    dm = c.get_data_manager_device_readwrite()
    names = ["density_dry", "uvel", "vvel", "wvel", "temp"] + c.get_tracer_names()
    if w.get("init") == "supercell":
        d.init_idealized_state_and_tracers(c, "supercell")
    else:
        for n in names:
            dm.get(n).copy_(torch.from_numpy(case["fields"][n]))
    d.declare_current_profile_as_hydrostatic(c)
    init = {n: dm.get(n).clone() for n in names}
    ncell = w["nz"] * w["ny"] * w["nx"] * nloc
    mass0 = d.compute_mass(c)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize(dev)

    def allmax(x):
        t = torch.tensor([x], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    def allsum(x):
        t = torch.tensor([x], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.SUM)
        return float(t.item())

    # ---- warm-up; the first step doubles as this rank's parity check: members are independent, so the first members of the
    # shard must equal what the CPU oracle gives for them alone (same sub-cycle count). The oracle is the checker here, never
    # the thing measured. ----
    ncycles = d.timeStep(c)
    nm = min(2, nloc)
    from oracle import oraclelib
    sub = {n: np.ascontiguousarray(init[n][..., :nm].cpu().numpy()) for n in names}
    o = oraclelib.OracleDycore(w["nz"], w["ny"], w["nx"], nm, case["xlen"], case["ylen"], case["zint"], tracers=case["tracers"],
                               consts=case["consts"], crm_dt=CRM_DT, ord=w["ord"])
    o.set_fields(sub)
    o.declare_hydrostatic()
    o.time_step(ncycles=ncycles)
    ref = o.get_fields()
    o.close()
    un = float(np.linalg.norm(ref["uvel"]))
    worst = 0.0
    for n in names:
        got = dm.get(n)[..., :nm].cpu().numpy()
        floor = 1e-3 * un if n in ("uvel", "vvel") else (1e-2 * case["consts"]["grav"] * CRM_DT * np.sqrt(got.size) if n == "wvel" else 1e-300)
        worst = max(worst, float(np.linalg.norm(got - ref[n]) / max(np.linalg.norm(ref[n]), floor)))
    parity = allmax(worst)
    for _ in range(max(warmup - 1, 0)):
        ncycles = d.timeStep(c)

    # ---- timed region: K steps, device-resident fields, CUDA events on the launching stream, no per-kernel events ----
    l0 = d.launch_count()
    sampler = ClockSampler(dev.index)
    if rank == 0:
        sampler.start()
        time.sleep(0.3)
    barrier()
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    ev0.record()
    cyc = []
    for _ in range(steps):
        cyc.append(d.timeStep(c))
    ev1.record()
    barrier()
    ms_total = allmax(ev0.elapsed_time(ev1))
    clocks = sampler.stop() if rank == 0 else None
    launches = d.launch_count() - l0
    upd = allsum(float(ncell) * sum(cyc))
    out = {"value": upd / (ms_total * 1e-3), "ms_per_step": ms_total / steps, "cyc": cyc, "launches": launches, "clocks": clocks,
           "ncell": ncell, "nloc": nloc, "names": names, "nt": len(case["tracers"]), "parity": parity, "updates_per_step": upd / steps,
           "fused_tracer": d.fused_tracer()}

    # ---- per-kernel-class profile in a SEPARATE pass (an event pair around every launch perturbs the headline slightly) ----
    if do_profile:
        d.profile_begin()
        pst = max(2, min(steps, 5))
        for _ in range(pst):
            d.timeStep(c)
        out["prof"], out["prof_steps"] = d.profile_end(), pst

    # global conservation diagnostic (the only collective on this path; untimed)
    mass1 = d.compute_mass(c)
    gm0, gm1 = global_mass(mass0, nens_global), global_mass(mass1, nens_global)
    nt = len(case["tracers"])
    out["mass_drift_rel"] = float(abs(gm1[nt] - gm0[nt]) / abs(gm0[nt]))

    # ---- end-to-end through the C ABI with HOST buffers (pinned), H2D + step + D2H inside the timed region, every step ----
    if do_e2e:
        host = [torch.empty(init[n].shape, dtype=torch.float64).pin_memory() for n in names]
        for hbuf, n in zip(host, names):
            hbuf.copy_(init[n])
        harr = [hbuf.numpy() for hbuf in host]
        if args.host_slabs > 0:
            d.set_host_slabs(args.host_slabs)
        for _ in range(2):
            d.time_step_host(harr, CRM_DT)
        barrier()
        ee0, ee1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        t0 = time.perf_counter()
        ee0.record()
        ecyc = [d.time_step_host(harr, CRM_DT) for _ in range(steps)]
        ee1.record()
        torch.cuda.synchronize(dev)
        wall = time.perf_counter() - t0
        # timed on the device like `value` (events on the launching stream; every call ends with the results in the host buffers
        # and the streams drained, so the host's clock reads the same to a fraction of a per cent: reported beside it)
        te = allmax(ee0.elapsed_time(ee1) * 1e-3)
        eu = allsum(float(ncell) * sum(ecyc))
        fb = 8 * ncell * len(names)
        e2e = {"value": eu / te, "unit": "cell-updates/s", "h2d_bytes_per_step": int(allsum(fb)), "d2h_bytes_per_step": int(allsum(fb)),
               "steps": steps, "ms_per_step": 1e3 * te / steps, "wall_ms_per_step": 1e3 * allmax(wall) / steps, "host_slabs": d.host_slabs(), "mispredictions": d.host_mispredictions(), "numa": numa,
               "note": "awfl_time_step_host: pinned host fields -> device, Dycore::timeStep, device -> host, every step; the ensemble "
                       "is pipelined in member slabs so that copies of one slab overlap the kernels of another"}
        if link:
            mine, lo, tot = link
            copy_ms = fb / (lo * 1e9) * 1e3
            floor_ms = max(copy_ms, out["ms_per_step"])
            e2e["roofline"] = {"bound": "host_link" if copy_ms > out["ms_per_step"] else "kernels",
                               "bytes_each_way_per_gpu": fb, "link_gbs_each_way_per_gpu": lo, "link_gbs_each_way_box": tot,
                               "min_copy_ms": copy_ms, "kernel_ms": out["ms_per_step"], "floor_ms": floor_ms,
                               "achieved_ms": e2e["ms_per_step"], "frac": floor_ms / e2e["ms_per_step"],
                               "note": "floor = max(bytes each way / measured pinned-copy rate with both directions busy on every rank at "
                                       "once, device-resident step time): copies and kernels overlap at best"}
        out["e2e"] = e2e
    d.finalize()
    del init, c, d
    torch.cuda.empty_cache()
    return out


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--workload", default="C2", choices=sorted(WORKLOADS))
    ap.add_argument("--ord", type=int, default=None)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-strong-c3", action="store_true", help="skip the strong-scaling line (C3, nens_global=1024)")
    ap.add_argument("--no-gpu-reference", action="store_true", help="skip the reference's own YAKL-CUDA kernels on this GPU (N=1 only)")
    ap.add_argument("--host-slabs", type=int, default=0, help="member slabs pipelined by the host-buffer entry point (0 = library default)")
    ap.add_argument("--nens", type=int, default=0, help="experiments: override the members per GPU of the workload")
    args = ap.parse_args()
    w = dict(WORKLOADS[args.workload])
    if args.nens > 0:
        w["desc"] = w["desc"].replace("nens=%d" % w["nens"], "nens=%d" % args.nens)
        w["nens"] = args.nens
    if args.ord:
        w["ord"] = args.ord
        w["desc"] = w["desc"].replace("ord=5", "ord=%d" % args.ord)
    if args.impl == "reference":
        run_reference_arm(args, w)
        return
    args.warmup = max(args.warmup, 3)

    import torch
    import torch.distributed as dist
    from pam_b200 import roofline

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise RuntimeError("bench.py needs a CUDA device: the AWFL path has no CPU fallback")
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    from pam_b200.parallel import bind_to_gpu_numa_node
    numa = bind_to_gpu_numa_node(local_rank) if world > 1 and not os.environ.get("AWFL_NO_NUMA_BIND") else None
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)

    link = None if args.no_e2e else host_link_rate(dev, world)
    nens_global = w["nens"] * world
    r = run_workload(w, args, rank, world, dev, nens_global, args.steps, args.warmup, not args.no_e2e, True, numa, link)

    # ---- strong scaling: BASELINE.json configs[2], 3-D 32x32x58 with nens_global = 1024 fixed, sharded over the N ranks ----
    strong = None
    if args.workload == "C2" and not args.no_strong_c3 and not args.nens and not args.ord:
        w3 = dict(WORKLOADS["C3"])
        ng3 = 1024
        w3["nens"] = ng3 // world
        w3["desc"] = "3-D 32x32x58, nens_global=1024 (%d/GPU), ord=5, NT=1" % (ng3 // world)
        s = run_workload(w3, args, rank, world, dev, ng3, max(3, min(args.steps, 10)), 3, not args.no_e2e, False, numa, link)
        strong = {"value": s["value"], "unit": "cell-updates/s", "ms_per_step": s["ms_per_step"], "nens_global": ng3, "nens_per_gpu": s["nloc"],
                  "n_gpus": world, "scaling": "strong", "ncycles": s["cyc"][-1], "steps": len(s["cyc"]), "workload": "C3: " + w3["desc"],
                  "parity_worst_rel_l2": s["parity"], "mass_drift_rel": s["mass_drift_rel"],
                  "e2e": s.get("e2e"),
                  "note": "BASELINE.json configs[2]: the total work is fixed, every rank owns 1024/N members; read value(N)/value(1) against "
                          "the >= 7x target at N = 8"}

    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return

    # ---- roofline of the dominant kernel (largest share of device time), measured live with CUDA events ----
    prof, cyc, ncell, nt, names = r["prof"], r["cyc"], r["ncell"], r["nt"], r["names"]
    tot_ms = sum(v[0] for v in prof.values())
    dom = max(prof, key=lambda k: prof[k][0])
    dom_ms, dom_n = prof[dom]
    ndir = 2 if w["ny"] == 1 else 3
    fp64_peak = measure_fp64_peak(dev, w)
    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except Exception:
        pass
    hbm_peak = peaks.get("hbm_gbs", 6650.0)
    hbm_src = "measured (MEASURED_PEAKS.json)" if "hbm_gbs" in peaks else "fallback (B200_PROFILING.md)"
    # 2-D: the x / z state march (bit 0 / 1) carries the first tracer; the tracer march of that direction runs for the others
    fused_bits = r.get("fused_tracer", 0)
    fused_dom = bool(fused_bits >> (1 if dom.endswith("_z") else 0) & 1) and not dom.endswith("_y")
    if dom.startswith("state_"):
        fl = roofline.state_sweep_flops_per_cell(w["ord"], sim2d=w["ny"] == 1, fused_tracer=fused_dom) * ncell
        by = roofline.state_sweep_bytes_per_cell(last=dom == "state_z", fused_tracer=fused_dom) * ncell
    elif dom.startswith("tracer_") and dom != "tracer_upd":
        fl = roofline.tracer_sweep_flops_per_cell(w["ord"]) * ncell * (nt - (1 if fused_dom else 0))
        by = 8 * 4 * ncell * (nt - (1 if fused_dom else 0))
    elif dom == "tracer_upd" and w["ny"] == 1:
        fl, by = 60.0 * ncell * nt, 8 * 5 * ncell * nt       # fused 2-D FCT + update: reads 2 fluxes + 2 states, writes 1 state
    else:
        fl, by = 40.0 * ncell * nt, 8 * 12 * ncell * nt
    traffic, executed = None, None
    try:
        tj = json.load(open(os.path.join(ROOT, "profiles", "traffic.json")))
        ent = tj.get(args.workload, {}).get(dom, {}) if w["ord"] == 5 else {}
        traffic = ent.get("bytes")
        if "fp64_pipe_pct" in ent:
            executed = {"fp64_pipe_frac": ent["fp64_pipe_pct"] / 100.0, "fp64_inst_per_cell": ent.get("fp64_inst_per_cell"),
                        "source": ent.get("source"),
                        "note": "what the hardware executed for the same launch (ncu): the kernel needs fewer FP64 instructions than the "
                                "reference algorithm has flops (row-form WENO, shared denominators), so `frac` (algorithmic) reads higher"}
    except Exception:
        pass
    # one "launch" of a class = its work of one Runge-Kutta stage (a vertical march that runs page by page -- ord 9 with a
    # paged table -- is several kernel launches per stage: their times add up)
    stage_launches = r["prof_steps"] * cyc[-1] * 3
    dur = dom_ms / max(stage_launches if dom not in ("in", "out") else dom_n, 1) * 1e-3
    roof = {"bound": "fp64", "kernel": dom, "achieved": fl / dur / 1e12, "peak": fp64_peak, "unit": "TFLOP/s",
            "frac": fl / dur / 1e12 / fp64_peak, "traffic": traffic, "executed": executed,
            "peak_source": "DFMA-chain microbenchmark run in this process (awfl_measure_fp64_peak); FP64 peak is not in MEASURED_PEAKS.json",
            "flops_per_launch": fl, "avg_launch_ms": dur * 1e3, "share_of_step": dom_ms / tot_ms,
            "hbm": {"achieved": by / dur / 1e9, "peak": hbm_peak, "unit": "GB/s", "frac": by / dur / 1e9 / hbm_peak,
                    "bytes_per_launch": by, "peak_source": hbm_src},
            "note": "the AWFL step is FP64-pipe bound (SURVEY.md 8d: 32-60 flop/B vs a ridge of ~6), so the binding roofline "
                    "is the FP64 one; `achieved` counts the REFERENCE algorithm's flops (pam_b200/roofline.py, frozen from the "
                    "instrumented count of tools/count_flops.py), launch time averaged over all launches of the class in a separate "
                    "profiling pass; `traffic` and `hbm.bytes_per_launch` describe a stage-2/3 launch of the class. In a 2-D workload the "
                    "state marches leave out the reconstruction of v, whose flux the reference computes and then discards "
                    "(Dycore.h:565): those flops are NOT counted in `achieved` (7 reconstructions per cell and direction, not 8); "
                    "where `fused_tracer` says so the state march also carries the first tracer (its reconstruction and flux are "
                    "counted in the march that executes them)",
            "fused_tracer": fused_bits}
    ms_total = r["ms_per_step"] * args.steps
    step_fl = roofline.flops_per_cell_update(w["ord"], w["nt"], ndir) * float(ncell) * sum(cyc)
    step_by = roofline.bytes_per_cell_update(w["nt"]) * float(ncell) * sum(cyc)
    whole = {"tflops": step_fl / (ms_total * 1e-3) / 1e12, "frac_fp64": step_fl / (ms_total * 1e-3) / 1e12 / fp64_peak,
             "gbs_algorithmic": step_by / (ms_total * 1e-3) / 1e9, "frac_hbm": step_by / (ms_total * 1e-3) / 1e9 / hbm_peak}
    ps = r["prof_steps"]
    kernels = {k: {"ms_per_step": v[0] / ps, "launches_per_step": v[1] / ps, "share": v[0] / tot_ms} for k, v in prof.items() if v[1]}

    cpu_baseline = None
    if not args.no_cpu_baseline and world == 1:
        try:
            cpu_baseline, _, _, _ = cpu_reference_rate(w)
        except Exception as ex:   # the oracle always exists in a built tree; say why if not
            cpu_baseline = {"value": None, "unit": "cell-updates/s", "cores": os.cpu_count(), "kind": "reference", "sample": "failed: %s" % ex}

    # ---- the reference's OWN kernels on this GPU (its YAKL CUDA backend compiled for sm_100a): a second, same-GPU baseline ----
    gpu_reference = None
    if world == 1 and not args.no_gpu_reference:
        # in a process of its own: it brings its own CUDA runtime instance and a YAKL pool of ~28 GB, and none of this repo's code
        try:
            from oracle import pamcuda
            if pamcuda.available("ref", w["ord"]):
                cmd = [sys.executable, os.path.join(ROOT, "tools", "gpu_reference_bench.py"), "--workload", args.workload, "--steps", "3",
                       "--ord", str(w["ord"]), "--nens", str(w["nens"])]
                p = subprocess.run(cmd, capture_output=True, text=True, timeout=600, env=dict(os.environ, CUDA_VISIBLE_DEVICES=str(local_rank)
                                   if "CUDA_VISIBLE_DEVICES" not in os.environ else os.environ["CUDA_VISIBLE_DEVICES"]))
                if p.returncode != 0:
                    raise RuntimeError("rc %d: %s" % (p.returncode, (p.stderr or p.stdout)[-300:]))
                gpu_reference = json.loads(p.stdout.strip().splitlines()[-1])
                gpu_reference["speedup_resident"] = r["value"] / gpu_reference["value"]
                if r.get("e2e"):
                    gpu_reference["speedup_e2e"] = r["e2e"]["value"] / gpu_reference["value"]
            else:
                gpu_reference = {"value": None, "unavailable": "oracle/_ref/libpam_ref_ord%d_cuda.so not built" % w["ord"]}
        except Exception as ex:
            gpu_reference = {"value": None, "unavailable": "failed: %s" % ex}

    line = {"metric": "AWFL dycore cell-updates/sec", "value": r["value"], "unit": "cell-updates/s", "n_gpus": world,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": r["ms_per_step"], "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": make_config(args, w, world, cyc[-1]),
            "e2e": r.get("e2e"), "gpu_launches": r["launches"], "clocks": r["clocks"], "roofline": roof, "whole_step": whole, "kernels": kernels,
            "cpu_baseline": cpu_baseline, "gpu_reference": gpu_reference, "strong_c3": strong,
            "parity_worst_rel_l2": r["parity"], "mass_drift_rel": r["mass_drift_rel"]}
    print(json.dumps(line))
    if world > 1:
        dist.destroy_process_group()


def measure_fp64_peak(dev, w):
    """DFMA-chain peak through a throw-away handle of a tiny configuration (the workload's own handle is already released)."""
    from pam_b200 import PamCoupler, Dycore
    c = PamCoupler(dev)
    c.set_option("crm_dt", 1.0)
    c.allocate_coupler_state(8, 1, 8, 32)
    c.set_grid(8000.0, 1000.0, np.arange(9) * 100.0)
    c.add_tracer("water_vapor", "", True, True)
    d = Dycore(5)
    d.init(c)
    v = d.measure_fp64_peak(1 << 14)
    d.finalize()
    return v


if __name__ == "__main__":
    main()
