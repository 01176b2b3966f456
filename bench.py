#!/usr/bin/env python
"""bench.py — FV3 dycore step throughput of the B200-native hot path (pace_b200) and of the CPU baseline.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--workload c384|c128|c48] [--impl reference]
    python bench.py --workload tracer_c192 | mapsingle_c384_l91_f32     (BASELINE.json configs 4 and 5, one GPU)
    python bench.py --workload coupling_c384                            (the driver's dycore <-> physics coupling calls)
    python bench.py --workload egress_c384                              (diagnostics stored every step vs no output)

One "step" = one DynamicalCore.step_dynamics (k_split=1, n_split=6: 6 acoustic substeps + tracer advection +
vertical remap) on synthetic inputs of the named shape (pace_b200/synthetic.py).  Metric: simulated-days/day =
dt_atmos / wall-seconds-per-step (NDSL/ndsl/performance/report.py:116-129 without the /365).
Prints ONE JSON line (rank 0).
"""
import argparse
import ctypes
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
if "reference" in sys.argv:  # before anything loads an OpenMP runtime: torchrun exports OMP_NUM_THREADS=1
    os.environ["OMP_NUM_THREADS"] = str(os.cpu_count() or 1)

WORKLOADS = {
    # name: (tile size, layout, description)
    "c384": (384, 2, "Baroclinic-shaped C384 L79 fp64, 6 tiles x 2x2 = 24 subdomains of 192x192x79, k_split=1 n_split=6"),
    "c128": (128, 1, "Baroclinic-shaped C128 L79 fp64, 6 subdomains of 128x128x79, k_split=1 n_split=6"),
    "c48": (48, 1, "Baroclinic-shaped C48 L79 fp64, 6 subdomains, k_split=1 n_split=6"),
    "c24": (24, 1, "Baroclinic-shaped C24 L79 fp64, 6 subdomains, k_split=1 n_split=6"),
}
DT_ATMOS = 225.0
N_SPLIT = 6
NPZ = 79
# algorithmic (compulsory) bytes per cell: SURVEY.md section 8(d) / DESIGN.md section 5
BYTES_PER_CELL_FVTP2D_INNER = 4 * 8  # inner PPM sweeps: read q, crx, cry ; write 2 fluxes -> counted 3R + 2W, see DESIGN


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        return json.load(open(p)), "measured"
    return {"hbm_gbs": 6650.0}, "fallback"


class ClockSampler(threading.Thread):
    def __init__(self, index=0):
        super().__init__(daemon=True)
        self.index = index
        self.samples = []
        self.stop_flag = False

    def run(self):
        q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
             "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")
        while not self.stop_flag:
            try:
                out = subprocess.run(["nvidia-smi", "-i", str(self.index), "--query-gpu=" + q, "--format=csv,noheader,nounits"],
                                     capture_output=True, text=True, timeout=5).stdout.strip()
                if out:
                    self.samples.append([x.strip() for x in out.split(",")])
            except Exception:
                pass
            time.sleep(0.2)

    def summary(self):
        if not self.samples:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["unavailable"]}
        sm = sorted(int(s[0]) for s in self.samples if s[0].isdigit())
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = [n for i, n in enumerate(names) if any(s[2 + i].lower().startswith("active") for s in self.samples)]
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": int(self.samples[0][1]), "reasons": reasons,
                "samples": len(self.samples)}


def set_host_threads(n):
    """Thread count of the OpenMP runtime the CPU port uses.  The environment variable only counts before the runtime
    is loaded (torch loads libgomp, and torchrun exports OMP_NUM_THREADS=1), so the runtime is told directly."""
    os.environ["OMP_NUM_THREADS"] = str(n)
    try:
        ctypes.CDLL("libgomp.so.1").omp_set_num_threads(int(n))
    except OSError:
        pass


def cpu_port_step_seconds(n_tile, layout, threads, steps=1, warmup=1):
    """Seconds per dycore step of the CPU port (oracle/cpu_port.py: the kernel sources compiled for the host with
    OpenMP, verified bit-exact against the reference's golden vectors) at C<n_tile>."""
    from oracle import cpu_port
    from pace_b200.synthetic import make_engine

    set_host_threads(threads)
    lib = cpu_port.load(openmp=True)
    eng = make_engine(n_tile, layout, npz=NPZ, device="cpu", _lib_handle=lib, n_split=N_SPLIT, k_split=1, dt_atmos=DT_ATMOS)
    for _ in range(warmup):
        eng.step_dynamics()
    t0 = time.time()
    for _ in range(steps):
        eng.step_dynamics()
    return (time.time() - t0) / steps


def host_sample_for(n_tile, layout):
    """The largest configuration of the CPU port that fits the host's free memory with a margin (x1.8 of an estimate
    that counts 150 fields; the unfused host build allocates ~140): the workload itself when it does — 92 GB at C384,
    i.e. a box with 166 GB free — else a quarter / sixteenth of its cells."""
    try:
        import psutil

        avail = psutil.virtual_memory().available
    except Exception:
        avail = 32e9
    for tile, lay in ((n_tile, layout), (n_tile // 2, 1), (n_tile // 4, 1), (48, 1)):
        n = tile // lay
        need = 150 * 6 * lay * lay * (NPZ + 1) * (n + 7) * (n + 8) * 8
        if tile >= 48 and need * 1.8 < avail:
            return tile, lay
    return 48, 1


REFERENCE_NUMPY = {  # the reference itself cannot run on the GPU box; measured in the build container (profiles/)
    "what": "UNMODIFIED reference, pyFV3 numpy backend, one step_dynamics at C48 L79 with n_split=1, 6 processes on 8 cores",
    "seconds_per_step": 55.97, "source": "profiles/r02_reference_numpy_c48.json",
    "gt_cpu_ifirst": "not buildable on this box (needs gridtools-cpp and boost; no network)",
}


def run_reference(args):
    """--impl reference: the CPU implementation of the path on the box's host cores.  The reference itself is
    Python (GT4Py numpy backend) with a compiled-stencil cache and cannot travel to the GPU box, so this arm times the
    CPU port (kind "port") with every host thread: the workload itself when the host has the memory for it, otherwise
    a bounded sample whose own time per step is printed next to the value scaled to the workload by cell count."""
    cores = os.cpu_count() or 1
    n_tile, layout, desc = WORKLOADS[args.workload]
    tile, lay = host_sample_for(n_tile, layout)
    same = (tile, lay) == (n_tile, layout)
    # one step of the full workload takes about a minute on 16 cores: one warm-up and one timed step
    k = 1 if same else max(1, min(args.steps, 2))
    dt = cpu_port_step_seconds(tile, lay, cores, steps=k, warmup=1)
    scale = (n_tile / tile) ** 2
    ms = dt * scale * 1000.0
    value = DT_ATMOS / (ms / 1000.0)
    sample = ("the workload itself: %d timed step(s) of C%d (%d subdomains) on the host after one warm-up step" % (k, tile, 6 * lay * lay)) if same else (
        "%d timed step(s) at C%d L79 (6 subdomains, %.0f ms each), scaled by cell count (x%.0f) to the workload"
        % (k, tile, dt * 1000.0, scale))
    line = {
        "impl": "reference", "metric": "dycore simulated-days/day", "value": value, "unit": "simulated days / day",
        "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms, "higher_is_better": True,
        "scaling": "strong", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": {"workload": desc},
        "steps_timed": k, "sample_ms_per_step": dt * 1000.0, "sample_is_workload": same,
        "cpu_baseline": {"value": value, "unit": "simulated days / day", "cores": cores, "kind": "port", "sample": sample},
        "reference_numpy_backend": REFERENCE_NUMPY,
        "e2e": {"value": value, "unit": "simulated days / day", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }
    print(json.dumps(line))


def side_workload(args):
    """BASELINE.json configs 4 and 5 as stand-alone bench lines (one GPU): `--workload tracer_c192` times
    TracerAdvection.__call__ (pace_b200_tracer_advection: 3 sub-cycles, 8 DENSE tracers, C192 L79 fp64, six tiles);
    `--workload mapsingle_c384_l91_f32` times MapSingle.__call__ in the reference's fp32 mode (pace_b200_map_single_f32,
    kord 9, iv 1, C384 L91, six tiles).  A step = one call; CUDA events; inputs >> L2."""
    import torch

    from pace_b200.native import NativeCore

    pk, pk_kind = peaks()
    dev = "cuda:0"
    torch.cuda.set_device(0)
    if args.workload == "coupling_c384":
        # the two coupling calls of the driver's time loop around its physics, C384 L79, 24 subdomains;
        # what the physics hands back is set once (perturbed copies of the state), outside the timing
        from pace_b200.engine import COUPLING_PHYSICS_OUT
        from pace_b200.synthetic import make_engine, setup_coupling

        n, nsub = 384, 6
        eng = make_engine(n, 2, npz=NPZ, device=dev, n_split=N_SPLIT, k_split=1, dt_atmos=DT_ATMOS)
        eng.step_dynamics()
        setup_coupling(eng, n, 2)
        core = eng.core
        eng.dycore_to_physics()
        t = core.t
        t["phy.prsi"][:, 0] = float(eng.cfg["ptop"])
        t["phy.prsi"][:, 1:] = float(eng.cfg["ptop"]) + torch.cumsum(t["delp"][:, :NPZ], dim=1)
        g = torch.Generator(device="cpu").manual_seed(0)
        for name in COUPLING_PHYSICS_OUT[1:]:
            src = {"physics_updated_specific_humidity": "qvapor"}.get(name, name.replace("physics_updated_", ""))
            r = torch.rand(t["phy." + src].shape[1:], generator=g, dtype=torch.float64).to(dev)
            if src in ("ua", "va"):
                t["phy." + name].copy_(t["phy." + src] + 1.5 * (r - 0.5))
            elif src == "pt":
                t["phy." + name].copy_(t["phy." + src] * (1.0 + 2.0e-3 * (r - 0.5)))
            else:
                t["phy." + name].copy_(t["phy." + src] * (1.0 + 0.05 * (r - 0.5)) + 1.0e-7 * r)

        def call():
            eng.dycore_to_physics()
            eng.update_atmos_state()

        cells = nsub * n * n * NPZ
        # field passes: 16 R + 16 W (copies), 24 R + 10 W (tendencies, tracers, delp), 3 (fill), 10 (moist_cv), 4
        # (pressures), 6 + 2 (D-grid winds, zeroing), 4 (A-grid winds) = 95
        bytes_per_cell, what, dtype = 95 * 8.0, "DycoreToPhysics + UpdateAtmosphereState, C384 L79, 24 subdomains", "f64"
        metric, unit = "coupling calls per second", "calls / s"
    elif args.workload == "tracer_c192":
        from pace_b200.synthetic import make_engine

        n, nsub = 192, 6
        eng = make_engine(n, 1, npz=NPZ, device=dev, n_split=N_SPLIT, k_split=1, dt_atmos=DT_ATMOS)
        eng.step_dynamics()  # leaves mass fluxes / Courant numbers of a real acoustic loop in mfxd, mfyd, cxd, cyd
        core = eng.core
        qv = core.t["qvapor"]
        for m, name in enumerate(eng_tracers()):
            if name != "qvapor":  # every species dense: a scaled copy of the vapour field with its own offset
                core.t[name].copy_(qv * (0.1 + 0.05 * m) + 1.0e-6 * (m + 1))
        call = lambda: eng.tracer_advection()  # noqa: E731
        cells = nsub * n * n * NPZ
        bytes_per_cell, what, dtype = 656.0, "TracerAdvection.__call__, 8 dense tracers, C192 L79, six tiles", "f64"
        metric, unit = "tracer advection calls per second", "calls / s"
    else:
        n, nsub, nz = 384, 6, 91
        core = NativeCore(n, n, nz, n_tiles=nsub, device=dev)
        g = torch.Generator(device="cpu").manual_seed(0)
        shape2 = (nsub, core.nj, core.sj)
        ps = (95000.0 + 8000.0 * torch.rand(shape2, generator=g)).to(dev)
        thick = (0.2 + 1.6 * torch.rand((nsub, nz) + shape2[1:], generator=g)).to(dev)
        thick = thick * ((ps - 300.0) / thick.sum(dim=1))[:, None]
        pe1 = torch.zeros((nsub, nz + 1) + shape2[1:], dtype=torch.float64, device=dev)
        pe1[:, 0] = 300.0
        pe1[:, 1:] = 300.0 + torch.cumsum(thick.double(), dim=1)
        eta = (torch.arange(nz + 1, dtype=torch.float64, device=dev) / nz) ** 1.3
        pe2 = 300.0 + (pe1[:, nz] - 300.0)[:, None] * eta[None, :, None, None]
        pe1, pe2 = pe1.float().contiguous(), pe2.float().contiguous()
        pe2[:, nz] = pe1[:, nz]
        zz = torch.linspace(0.0, 1.0, nz + 1, device=dev)[None, :, None, None]
        q0 = (250.0 + 40.0 * zz + 1.5 * torch.randn(pe1.shape, generator=g).to(dev)).float().contiguous()
        q = q0.clone()
        del thick, ps

        def call():
            q.copy_(q0)  # the remap is in place: every call starts from the same field (the copy is inside the timing)
            core.map_single_f32(q, pe1, pe2, None, iv=1, kord=9, qmin=0.0)

        cells = nsub * n * n * nz
        # q read + written, pe1 and pe2 read (fp32), + the restoring copy (read + write)
        bytes_per_cell, what, dtype = 6 * 4.0, "MapSingle.__call__ fp32 (kord 9, iv 1), C384 L91, six tiles", "f32"
        metric, unit = "MapSingle calls per second", "calls / s"
    lib = core.lib
    for _ in range(max(args.warmup, 3)):
        call()
    torch.cuda.synchronize()
    sampler = ClockSampler(0)
    sampler.start()
    l0 = lib.pace_b200_launch_count()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(args.steps):
        call()
    e1.record()
    torch.cuda.synchronize()
    sampler.stop_flag = True
    sampler.join(timeout=2)
    ms = e0.elapsed_time(e1) / args.steps
    achieved = cells * bytes_per_cell / (ms / 1000.0) / 1e9
    line = {
        "metric": metric, "value": 1000.0 / ms, "unit": unit, "n_gpus": 1, "steps": args.steps, "warmup": max(args.warmup, 3),
        "ms_per_step": ms, "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": dtype,
        "data": "synthetic", "config": {"workload": what, "cells": cells, "l2": "inputs >> L2, no flush needed"},
        "gpu_launches": int((lib.pace_b200_launch_count() - l0) // args.steps), "clocks": sampler.summary(),
        "roofline": {"bound": "hbm", "achieved": achieved, "peak": pk["hbm_gbs"], "peak_source": pk_kind, "unit": "GB/s",
                     "frac": achieved / pk["hbm_gbs"], "traffic": None, "algorithmic_bytes_per_step": cells * bytes_per_cell,
                     "bytes_per_cell": bytes_per_cell},
    }
    print(json.dumps(line))


def egress_workload(args):
    """the headline workload with diagnostics stored EVERY step (pace_b200.egress.MonitorDiagnostics: ten
    3-D fields, two column integrals, three level selections) against the same steps without output: what an output
    step costs the time loop.  The monitor only samples what it receives (a strided checksum per array): zarr / netCDF4
    are not in this image, and the cost of a file writer is the writer's, not the egress path's; a sink that is slower
    than the time loop shows up as back-pressure (wall time per step above the device time)."""
    import numpy as np
    import torch

    from pace_b200.egress import MonitorDiagnostics
    from pace_b200.synthetic import make_engine

    dev = "cuda:0"
    torch.cuda.set_device(0)
    eng = make_engine(384, 2, npz=NPZ, device=dev, n_split=N_SPLIT, k_split=1, dt_atmos=DT_ATMOS)
    names = ["u", "v", "w", "pt", "delp", "delz", "qvapor", "qliquid", "ua", "va"]

    class Monitor:
        def __init__(self):
            self.checksum, self.n = 0.0, 0

        def store(self, state):
            for k, v in state.items():
                if k == "time":
                    continue
                for q in v:
                    self.checksum += float(np.sum(q.data[::7, ::7]))
            self.n += 1

        def store_constant(self, state):
            pass

        def cleanup(self):
            pass

    mon = Monitor()
    diag = MonitorDiagnostics(mon, names, ["column_integrated_qvapor", "column_integrated_qliquid"],
                              [(40, ["pt", "w"]), (78, ["ua"])], dycore=eng)
    for _ in range(max(args.warmup, 3)):
        eng.step_dynamics()
        diag.store(0.0)
    torch.cuda.synchronize()
    sampler = ClockSampler(0)
    sampler.start()

    def timed(with_store):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        torch.cuda.synchronize()
        t0 = time.time()
        e0.record()
        for n in range(args.steps):
            eng.step_dynamics()
            if with_store:
                diag.store(float(n))
        e1.record()
        torch.cuda.synchronize()
        if with_store:
            diag._egress.cleanup()   # every stored step has reached the monitor
        return e0.elapsed_time(e1) / args.steps, (time.time() - t0) / args.steps * 1000.0

    plain_ms, _ = timed(False)
    l0 = eng.core.lib.pace_b200_launch_count()
    dev_ms, wall_ms = timed(True)
    launches = (eng.core.lib.pace_b200_launch_count() - l0) // args.steps
    sampler.stop_flag = True
    sampler.join(timeout=2)
    nbytes = diag._egress.bytes_per_store
    stored = mon.n
    diag.cleanup()
    line = {
        "metric": "dycore simulated-days/day with diagnostics stored every step", "value": DT_ATMOS / (wall_ms / 1000.0),
        "unit": "simulated days / day", "n_gpus": 1, "steps": args.steps, "warmup": max(args.warmup, 3),
        "ms_per_step": wall_ms, "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f64",
        "data": "synthetic",
        "config": {"workload": "C384 L79 fp64, 24 subdomains, MonitorDiagnostics.store after every step: %d 3-D fields, "
                               "2 column integrals, 3 level selections" % len(names)},
        "gpu_launches": int(launches), "clocks": sampler.summary(),
        "egress": {"ms_per_step_device": dev_ms, "ms_per_step_wall_until_the_monitor_has_everything": wall_ms,
                   "ms_per_step_without_output": plain_ms, "cost_of_an_output_step_ms": dev_ms - plain_ms,
                   "d2h_bytes_per_store": nbytes, "d2h_gbs_over_the_step": nbytes / (wall_ms / 1000.0) / 1e9,
                   "stores_received_by_the_monitor": stored},
    }
    print(json.dumps(line))


def eng_tracers():
    return ["qvapor", "qliquid", "qrain", "qice", "qsnow", "qgraupel", "qo3mr", "qsgs_tke"]


def state_hash(eng, dist, world):
    """SHA-256 over the prognostic fields (compute domain, subdomains in cube order) after the timed steps: the same
    number of steps from the same state gives the same hash at every GPU count when the partitioning changes nothing."""
    import hashlib

    import numpy as np

    core = eng.core
    n = core.nx
    digests = []
    arrs = {name: core.t[name].cpu().numpy() for name in ("u", "v", "w", "delp", "pt", "delz", "qvapor")}
    for t in range(core.nt):
        h = hashlib.sha256()
        for name, a in arrs.items():
            h.update(np.ascontiguousarray(a[t, :NPZ, 3:3 + n, 3:3 + n]).tobytes())
        digests.append(h.hexdigest())
    if dist is not None:
        allp = [None] * world
        dist.all_gather_object(allp, digests)
        digests = [d for part in allp for d in part]
    return hashlib.sha256("".join(digests).encode()).hexdigest()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--workload", default=None)
    ap.add_argument("--impl", default="pace_b200")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--e2e-full-state", action="store_true", help="e2e downloads all 32 DycoreState fields every step")
    ap.add_argument("--no-dense", action="store_true", help="skip the extra run with PACE_B200_ZERO_PATHS=0")
    args = ap.parse_args()
    if args.workload is None:
        args.workload = "c384"
    if args.workload == "egress_c384":
        return egress_workload(args)
    if args.workload in ("tracer_c192", "mapsingle_c384_l91_f32", "coupling_c384"):
        if int(os.environ.get("RANK", "0")) == 0:
            side_workload(args)
        return
    if args.impl == "reference":
        if int(os.environ.get("RANK", "0")) != 0:
            return  # under torchrun rank 0 alone times the CPU implementation
        return run_reference(args)

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))

    # The kernels skip work on data that is zero (rows / levels / columns / tiles of zeros; exact).  The baroclinic case
    # has seven of its eight tracer species and q_con identically zero, so the line also says what the same step costs
    # with those shortcuts switched off (a child process, before this one allocates its state).
    dense = None
    if world == 1 and not args.no_dense and os.environ.get("PACE_B200_ZERO_PATHS", "1") != "0":
        env = dict(os.environ, PACE_B200_ZERO_PATHS="0")
        cmd = [sys.executable, os.path.abspath(__file__), "--workload", args.workload, "--steps", str(min(args.steps, 3)),
               "--warmup", str(min(args.warmup, 2)), "--no-cpu-baseline", "--no-e2e", "--no-dense"]
        try:
            out = subprocess.run(cmd, env=env, capture_output=True, text=True, timeout=600).stdout.strip().splitlines()
            d = json.loads(out[-1])
            dense = {"ms_per_step": d["ms_per_step"], "value": d["value"], "unit": d["unit"],
                     "what": "same workload with PACE_B200_ZERO_PATHS=0 (no shortcuts for zero-valued tracers / q_con)"}
        except Exception as exc:  # the headline measurement does not depend on it
            dense = {"error": str(exc)[:200]}

    import torch
    from pace_b200 import _lib
    from pace_b200.synthetic import make_engine

    torch.cuda.set_device(local_rank)
    comm = None
    dist = None
    if world > 1:
        # one process per GPU; each holds a contiguous block of subdomains (halo transport: see config.placement)
        import torch.distributed as dist

        from pace_b200.distributed import TorchComm

        dist.init_process_group("nccl", device_id=torch.device("cuda:%d" % local_rank))
        comm = TorchComm()
    n_tile, layout, desc = WORKLOADS[args.workload]
    eng = make_engine(n_tile, layout, npz=NPZ, device="cuda:%d" % local_rank, comm=comm, n_split=N_SPLIT, k_split=1,
                      dt_atmos=DT_ATMOS)

    def fence():
        torch.cuda.synchronize()
        if dist is not None:
            dist.barrier()
            torch.cuda.synchronize()

    def max_over_ranks(x):
        if dist is None:
            return x
        t = torch.tensor([x], dtype=torch.float64, device="cuda:%d" % local_rank)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())
    core = eng.core
    lib = core.lib
    n_sub = 6 * layout * layout
    n = n_tile // layout
    cells = n_sub * n * n * NPZ

    for _ in range(args.warmup):
        eng.step_dynamics()
    fence()

    sampler = ClockSampler(local_rank)
    sampler.start()
    lib.pace_b200_profile_enable(1)
    l0 = lib.pace_b200_launch_count()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    fence()
    torch.cuda.profiler.start()  # lets `ncu --profile-from-start off` capture exactly the timed steps
    e0.record()
    for _ in range(args.steps):
        eng.step_dynamics()
    e1.record()
    fence()
    torch.cuda.profiler.stop()
    ms = max_over_ranks(e0.elapsed_time(e1) / args.steps)
    state_sha256 = state_hash(eng, dist, world)
    launches = (lib.pace_b200_launch_count() - l0) // args.steps
    lib.pace_b200_profile_enable(0)
    buf = ctypes.create_string_buffer(65536)
    lib.pace_b200_profile_report(buf, 65536)
    prof = {}
    for ln in buf.value.decode().splitlines():
        label, cnt, tot = ln.split()
        prof[label] = (int(cnt), float(tot))
    sampler.stop_flag = True
    sampler.join(timeout=2)
    value = DT_ATMOS / (ms / 1000.0)

    # roofline of the dominant labelled kernel
    pk, pk_kind = peaks()
    roof = None
    if prof:
        top = max(prof, key=lambda k: prof[k][1])
        cnt, tot = prof[top]
        per_launch_ms = tot / cnt
        bytes_per_cell = {"fvtp2d_inner_ppm": 5 * 8, "fvtp2d_qi_qj": 7 * 8, "fvtp2d_outer_ppm": 10 * 8,
                          "fvtp2d_fused": 8.4 * 8,
                          # dense launches of a step: delp, vorticity, zh (7 passes), w, pt (9 passes) per substep -> 7.8
                          "fvtp2d_march": 7.8 * 8, "fvtp2d_march_sparse": 9 * 8, "riem_solver_c": 8 * 8, "riem_solver3": 15 * 8, "remap_scalars_column": 40 * 8,
                          # c_sw: u, v, delp, pt, w read; uc, vc, ua, va, ut, vt, divgd, omga, delpc, ptc written
                          "c_sw_fused": 15 * 8,
                          # DelnFlux: q (+ mass) read, fx / fy read-modify-written or fx2 / fy2 / d2 written
                          "delnflux_march": 5.5 * 8, "delnflux_fused": 5.5 * 8,
                          # divergence damping loop: divg_d read; delpc, d_N, uc, vc written
                          "dd_march": 5 * 8}.get(top, 8 * 8)
        alg_bytes = cells / world * bytes_per_cell  # this GPU's share of the cells
        achieved = alg_bytes / (per_launch_ms / 1000.0) / 1e9
        # measured DRAM bytes per launch of that kernel from the committed ncu capture (profiles/traffic_r0*.json),
        # scaled by this GPU's share of the cells; null when no capture exists for this workload / kernel
        traffic = None
        for tname in ("traffic_r02.json", "traffic_r01.json"):  # the newest capture that has this kernel
            tpath = os.path.join(ROOT, "profiles", tname)
            rec = json.load(open(tpath)).get(args.workload, {}).get(top) if os.path.exists(tpath) else None
            if rec:
                traffic = rec["dram_bytes_per_launch"] / world
                break
        roof = {"kernel": top, "bound": "hbm", "achieved": achieved, "peak": pk["hbm_gbs"], "peak_source": pk_kind,
                "unit": "GB/s", "frac": achieved / pk["hbm_gbs"], "traffic": traffic,
                "algorithmic_bytes_per_launch": alg_bytes, "launches_timed": cnt,
                "ms_per_launch": per_launch_ms, "share_of_step": tot / args.steps / ms,
                "labelled_kernels_ms_per_step": {k: v[1] / args.steps for k, v in sorted(prof.items())}}

    # the sections the reference's own Timer brackets (fv_dynamics.py:551-590), on one extra step outside the timed
    # region, clocked with CUDA events on the launching stream; the acoustic substep's algorithmic GB/s is the second
    # half of BASELINE.json's metric: 97 field passes x 8 B = 776 B per cell per substep
    class EventTimer:
        def __init__(self):
            self.ev = {}

        def start(self, name):
            e = torch.cuda.Event(enable_timing=True)
            e.record()
            self.ev.setdefault(name, []).append([e, None])

        def stop(self, name):
            e = torch.cuda.Event(enable_timing=True)
            e.record()
            self.ev[name][-1][1] = e

        def ms(self):
            return {k: sum(a.elapsed_time(b) for a, b in v) for k, v in self.ev.items()}

    et = EventTimer()
    eng.step_dynamics(timer=et)
    fence()
    sect = {k: max_over_ranks(v) for k, v in sorted(et.ms().items())}
    sub_ms = sect["DynCore"] / N_SPLIT
    sub_bytes = cells / world * 776.0
    acoustic = {"ms_per_substep": sub_ms, "algorithmic_bytes_per_substep": sub_bytes, "bytes_per_cell": 776,
                "achieved": sub_bytes / (sub_ms / 1000.0) / 1e9, "unit": "GB/s", "peak": pk["hbm_gbs"],
                "frac": sub_bytes / (sub_ms / 1000.0) / 1e9 / pk["hbm_gbs"], "sections_ms": sect}

    # end-to-end through the public API for a host-resident model (pace_b200.engine.HostPipeline): every step uploads
    # the full DycoreState from pinned host memory and downloads the updated state; uploads / downloads of
    # neighbouring steps overlap the compute on their own streams.  Pipeline fill and drain are inside the timing.
    e2e = None
    if not args.no_e2e:
        from pace_b200.engine import PHYSICS_FIELDS, HostPipeline

        # uploads the fields a step reads, downloads what the reference's driver reads back after a step
        # (engine.PHYSICS_FIELDS: prognostics, A-grid winds, pe / peln, omga, ps, tracers); --e2e-full-state: all 32
        pipe = HostPipeline(eng, download_names=None if args.e2e_full_state else PHYSICS_FIELDS)
        hin, hout = pipe.pinned_like(), pipe.pinned_like()
        for k in pipe.names:
            hin[k].copy_(core.t[k])
        # pipeline fill (one upload) and drain (one download) are inside the timing: about one step's worth of copies,
        # spread over k_e2e steps
        k_e2e = min(max(args.steps, 3), 24)
        fence()
        t0 = time.time()
        pipe.run(k_e2e, hin, hout)
        fence()
        e2e_ms = max_over_ranks((time.time() - t0) / k_e2e * 1000.0)
        nbytes = pipe.bytes_per_step
        e2e = {"value": DT_ATMOS / (e2e_ms / 1000.0), "unit": "simulated days / day", "ms_per_step": e2e_ms,
               "h2d_bytes_per_step": pipe.upload_bytes_per_step * world, "d2h_bytes_per_step": nbytes * world,
               "steps": k_e2e, "fields": len(pipe.names), "fields_uploaded": len(pipe.upload_names),
               "fields_downloaded": len(pipe.download_names),
               "downloaded": "engine.PHYSICS_FIELDS (what the reference's driver and physics hand-over read after a step)" if not args.e2e_full_state else "the whole DycoreState",
               "h2d_gbs": pipe.upload_bytes_per_step / (e2e_ms / 1000.0) / 1e9, "d2h_gbs": nbytes / (e2e_ms / 1000.0) / 1e9,
               "h2d_copy_gbs": pipe.upload_bytes_per_step / (pipe.copy_ms[0] / 1000.0) / 1e9,
               "d2h_copy_gbs": nbytes / (pipe.copy_ms[1] / 1000.0) / 1e9,
               "fill_drain_ms": pipe.copy_ms[0] + pipe.copy_ms[1], "device_state_copies": pipe.nsets,
               "overlap": "upload(n+1) / download(n-1) on side streams during step n; three device copies of the state, "
                          "so an upload never waits for the previous download; h2d_gbs / d2h_gbs = bytes over the step time, "
                          "*_copy_gbs = bytes over the duration of the copies themselves (last step)"}
        del hin, hout, pipe

    cpu = None
    if not args.no_cpu_baseline and world == 1:
        cores = os.cpu_count() or 1
        sample_tile = 96 if n_tile >= 96 else 48
        dt = cpu_port_step_seconds(sample_tile, 1, cores, steps=1, warmup=1)
        scale = (n_tile / sample_tile) ** 2
        cpu = {"value": DT_ATMOS / (dt * scale), "unit": "simulated days / day", "cores": cores, "kind": "port",
               "sample": "one dycore step at C%d L79 (%.1f s on the host), scaled by cell count (x%.0f)" % (sample_tile, dt, scale),
               "reference_numpy_backend": REFERENCE_NUMPY}

    line = {
        "metric": "dycore simulated-days/day", "value": value, "unit": "simulated days / day", "n_gpus": world,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms, "higher_is_better": True, "scaling": "strong",
        "vs_baseline": None, "dtype": "f64", "data": "synthetic (C12 reference fixture resampled, see pace_b200/synthetic.py)",
        "config": {"workload": desc, "cells": cells, "placement": ("%d subdomains per GPU, halo exchange: %s" % (n_sub // world, "peer stores into symmetric memory over NVLink + device-side barrier" if eng.arena is not None else "torch.distributed isend / irecv")) if world > 1 else "%d subdomains on one GPU" % n_sub, "l2": "working set (%.0f GB) >> L2, no flush needed" % (cells * 8 * 130 / 1e9)},
        "state_sha256": state_sha256, "gpu_launches": int(launches), "clocks": sampler.summary(), "roofline": roof, "acoustic_substep": acoustic, "dense_path": dense,
        "cpu_baseline": cpu, "e2e": e2e,
    }
    if rank == 0:
        print(json.dumps(line))
    if dist is not None:
        dist.barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
