#!/usr/bin/env python
"""Benchmark of the FDTD time-stepping path (BASELINE.json metric: Yee cell-updates/s in Gcells/s).

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl b200|reference] [--config 1|2|3|app|thin] [--lattice nx,ny,nz]

A "step" is one full leapfrog step (H half-step then E half-step, six components) of the whole lattice.
--config 3 (default, the headline): BASELINE configs[2] — the 512^3 heterogeneous lattice inside the lossy absorbing
        layer.  N > 1: the same kind of lattice grown so every GPU keeps 512^3 cells (weak scaling),
        z-slab decomposed: N=2 1024x512x512, N=4 1024x1024x512, N=8 1024^3; halo planes are stored into the neighbours'
        ghost planes over NVLink peer memory with device-side flags (--transport p2p, default) or exchanged with NCCL
        send/recv (--transport nccl).  Before the timed region an N>1 run checks that a slab-decomposed 256x256x(64N)
        lattice run through the same transport is bit-identical to the single-domain run (`slab_parity`).
--config 1 | 2 | app | thin: the other lattices the reference runs — 64^3 vacuum (configs[0]), 256^3 inclusions + probes
        (configs[1]), the application's own 100x100x1 test lattice (cem-app/src/composer/mod.rs:841-869), 4096x4096x1 —
        one GPU, same JSON line, CPU leg on the SAME lattice.
Timing: W warm-up steps, then P >= 5 passes of EXACTLY K steps each, every pass bracketed by barrier + synchronize and timed
with CUDA events on the launching stream; `value` is the MEDIAN pass (max over ranks per pass).  The end-to-end passes (`e2e`:
host buffers, per-step H2D forcing and D2H probe row inside the timed region) alternate with the device-resident ones, so both
see the same clocks; clocks are sampled over all of them.
`--impl reference` times the CPU restatement of the reference's own CPU backend (oracle/, f64 AoS, every host core) — on the
whole lattice at N=1 (same config as the GPU arm), on a bounded z-slab sample of the N-GPU lattice for N>1; the Rust reference
itself cannot be built here.  Prints ONE JSON line on rank 0.
"""
from __future__ import annotations

import argparse
import hashlib
import json
import math
import os
import statistics
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

BYTES_PER_CELL_HALF_STEP = 37.0   # 6 fp32 reads + 3 fp32 writes + 1 B material index
BYTES_PER_CELL_STEP = 74.0
BYTES_PER_CELL_FUSED = 49.0       # fused H+E pass: 6 fp32 reads + 6 fp32 writes + 1 B index
FALLBACK_HBM_GBS = 6650.0         # /opt/skills/guides/B200_PROFILING.md fallback
CPU_BYTES_PER_CELL = 136 + 4      # oracle state (cpu/mod.rs:217-221: 32 coeff + 4x24 field + 8 source idx) + the u32 id lattice


def lattice_for(n_gpus: int) -> tuple:
    return {1: (512, 512, 512), 2: (1024, 512, 512), 4: (1024, 1024, 512), 8: (1024, 1024, 1024)}.get(
        n_gpus, (512, 512, 512 * n_gpus))


def host_cores() -> int:
    """Every core this process may run on — NOT OMP_NUM_THREADS, which torch.distributed.run exports as 1
    (≙ FdtdCpuBackend::multi_threaded(None) = all cores, cem-solver/src/fdtd/cpu/mod.rs:148-157)."""
    try:
        return max(len(os.sched_getaffinity(0)), 1)
    except AttributeError:
        return max(os.cpu_count() or 1, 1)


def mem_available_bytes() -> int:
    try:
        with open("/proc/meminfo") as f:
            for line in f:
                if line.startswith("MemAvailable:"):
                    return int(line.split()[1]) * 1024
    except OSError:
        pass
    return 0


def read_peak_hbm() -> tuple:
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    try:
        with open(path) as f:
            return float(json.load(f)["hbm_gbs"]), "measured (MEASURED_PEAKS.json)"
    except Exception:
        return FALLBACK_HBM_GBS, "fallback (B200_PROFILING.md)"


class ClockSampler:
    """Samples SM clocks and throttle reasons with nvidia-smi while the timed region runs."""

    FIELDS = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
              "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")
    NAMES = ("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap")

    def __init__(self, gpu_index: int):
        self.rows = []
        self.proc = None
        self.gpu = gpu_index

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.FIELDS}", "--format=csv,noheader,nounits", "-lms", "20", "-i", str(self.gpu)],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.thread = threading.Thread(target=self._read, daemon=True)
            self.thread.start()
        except OSError:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append((time.monotonic(), line.strip()))

    def stop(self):
        if self.proc is not None:
            self.proc.terminate()
            try:
                self.proc.wait(timeout=5)
            except subprocess.TimeoutExpired:
                self.proc.kill()

    def count(self, t0: float) -> int:
        return sum(1 for t, _ in self.rows if t >= t0)

    def summary(self, t0: float, t1: float):
        sm, sm_max, reasons = [], [], set()
        for t, line in self.rows:
            if t < t0 or t > t1:
                continue
            parts = [p.strip() for p in line.split(",")]
            if len(parts) < 7:
                continue
            try:
                sm.append(float(parts[0])); sm_max.append(float(parts[1]))
            except ValueError:
                continue
            for name, val in zip(self.NAMES, parts[3:7]):
                if val.lower().startswith("active"):
                    reasons.add(name)
        if not sm:
            return None
        return {"sm_mhz": statistics.median(sm), "sm_max_mhz": max(sm_max), "reasons": sorted(reasons), "samples": len(sm)}


# ---- workloads ---------------------------------------------------------------------------------------------------------

def build_workload(lattice, z_begin, z_end):
    """Planes [z_begin, z_end) of the config-3 family: heterogeneous lattice in the graded lossy layer."""
    from fdtd_b200 import configs
    from fdtd_b200.solver import GaussianPulse

    pal = configs.heterogeneous_palette()
    ids = configs.heterogeneous_slab_labels(lattice, z_begin, z_end)
    nx, ny, nz = lattice
    src = GaussianPulse(12.5, 2.5).with_amplitudes((0.0, 0.0, 50.0), (0.0, 0.0, 0.0))
    sources = [((nx // 2, ny // 2, nz // 2), src)]
    probes = [(min(nx // 2 + 16 * (k + 1), nx - 1), ny // 2, nz // 2) for k in range(4)] + [(nx // 2, min(ny // 2 + 16 * (k + 1), ny - 1), nz // 2) for k in range(4)]
    return configs.Workload(f"heterogeneous{nx}x{ny}x{nz}", tuple(lattice), configs.REDUCED_RES, configs.PhysicalConstants.REDUCED,
                            pal, ids, sources, probes, 0)


def case_workload(case: str, lattice=None):
    """(workload, description) of a single-GPU bench case other than the headline."""
    from fdtd_b200 import configs
    from fdtd_b200.solver import GaussianPulse, Material, PhysicalConstants, Resolution

    if case == "1":
        w = configs.config1_vacuum64()
        c = w.lattice[0] // 2
        w.probes = [(c + 8, c, c), (c, c + 8, c), (c, c, c + 8)]
        return w, "64x64x64 vacuum cube, Gaussian-pulse point source, 3 probes (BASELINE configs[0])"
    if case == "2":
        w = configs.config2_inclusions()
        return w, "256x256x256 lattice with dielectric / lossy / magnetic inclusions and a conductor slab, CW point source, 8 probes (BASELINE configs[1])"
    if case in ("app", "thin"):
        # the application's canned solver config (cem-app/src/composer/mod.rs:841-869): volume 1x1x0 at dx = 0.01 -> 100x100x1,
        # dt = 0.001, REDUCED constants, default material sigma = 10 "to introduce dissipation"; one Gaussian source in the middle
        n = 100 if case == "app" else 4096
        lat = tuple(lattice) if lattice else (n, n, 1)
        res = Resolution((0.01, 0.01, 0.01), 0.001)
        src = GaussianPulse(0.05, 0.01).with_amplitudes((0.0, 0.0, 50.0), (0.0, 0.0, 0.0))
        c = (lat[0] // 2, lat[1] // 2, lat[2] // 2)
        probes = [(min(c[0] + 10, lat[0] - 1), c[1], c[2]), (c[0], min(c[1] + 10, lat[1] - 1), c[2])]
        w = configs.Workload(f"app{lat[0]}x{lat[1]}x{lat[2]}", lat, res, PhysicalConstants.REDUCED, [Material(eletrical_conductivity=10.0)], None,
                             [(c, src)], probes, 0)
        what = "the application's own test lattice" if case == "app" else "the application's test configuration on a large 2-D lattice"
        return w, f"{lat[0]}x{lat[1]}x{lat[2]} uniform sigma=10 medium, dx=0.01, dt=0.001, Gaussian point source, 2 probes ({what}, cem-app/src/composer/mod.rs:841-869)"
    raise SystemExit(f"unknown --config {case}")


def workload_config(args, lattice, n_gpus: int, description=None) -> dict:
    """The `config` object of the JSON line: what is computed, identical for the GPU arm and the reference arm."""
    nx, ny, nz = lattice
    cells = nx * ny * nz
    if description is None:
        description = (f"heterogeneous {nx}x{ny}x{nz} lattice (64 balls of 4 material kinds) inside the 16-cell graded lossy layer, "
                       "Gaussian point source, 8 probes (BASELINE configs[2]; "
                       + ("the same lattice for every N: strong scaling, BASELINE configs[3])" if args.scaling == "strong" else "weak-scaled 512^3 cells per GPU for N>1)"))
    state_mb = 25 * cells / n_gpus / 1e6  # six f32 fields + a one-byte index, one copy
    if state_mb < 126:
        l2 = (f"lattice state ({state_mb:.1f} MB per GPU) fits the 126 MB L2 and stays there on purpose: steps run back to back on resident "
              "data, as the application runs them; no flush")
    elif state_mb < 4 * 126:
        l2 = (f"no L2 flush between steps: the lattice state ({state_mb:.0f} MB per GPU, read once and written once per step) is "
              f"{state_mb / 126:.1f}x the 126 MB L2, so each step streams it from HBM (ncu dram bytes per launch confirm: profiles/)")
    else:
        l2 = f"no L2 flush between steps: the step's working set (>= 25 B/cell x {cells // n_gpus / 1e6:.0f}M cells per GPU) is far larger than the 126 MB L2"
    return {"workload": description, "lattice": [nx, ny, nz], "cells_per_gpu": cells // n_gpus, "decomposition": f"z-slabs x{n_gpus}", "l2": l2}


def cpu_sample(lattice, planes: int):
    """A z-slab sample of the benchmark lattice for the CPU legs: `planes` planes around the source plane."""
    nx, ny, nz = lattice
    planes = min(planes, nz)
    z0 = max(nz // 2 - planes // 2, 0)
    z1 = min(z0 + planes, nz)
    return z0, z1


def time_cpu_reference(lattice, planes: int, steps: int, warmup: int, threads=None, workload=None) -> dict:
    """Times the oracle's f64 restatement of FdtdCpuBackend (multi-threaded ≙ cpu/mod.rs:148-157).  `workload` given: that
    whole lattice.  Else planes [z0, z1) of the config-3 family lattice (planes >= nz: the whole lattice)."""
    import oracle
    from fdtd_b200 import configs

    oracle.build()
    cores = host_cores() if threads is None else int(threads)
    oracle.set_threads(cores, fast=True)
    nx, ny, nz = lattice
    if workload is not None:
        w, z0, z1 = workload, 0, nz
        pal = w.palette_rows()
        ids = w.ids
        pts = w.source_points()
        sub = lattice
    else:
        z0, z1 = cpu_sample(lattice, planes)
        sub = (nx, ny, z1 - z0)
        ids = configs.heterogeneous_slab_labels(lattice, z0, z1)
        pal = np.array([m.as_row() for m in configs.heterogeneous_palette()], dtype=np.float64)
        w = build_workload(lattice, z0, z1)
        pts = [(nx // 2, ny // 2, nz // 2 - z0)]
    s = oracle.RefSolver(sub, w.resolution.spatial, w.resolution.temporal, precision=oracle.F64, fast=True)
    s.set_materials(pal, ids, w.constants.vacuum_permittivity, w.constants.vacuum_permeability)
    tab = w.source_table(warmup + steps)
    s.run(warmup, pts, tab[:warmup])
    _, secs = s.run(steps, pts, tab[warmup:])
    cells = nx * ny * (z1 - z0)
    whole = (z1 - z0) == nz
    where = f"the whole {nx}x{ny}x{nz} lattice" if whole else f"planes [{z0},{z1}) of the {nx}x{ny}x{nz} lattice"
    return {"value": cells * steps / secs / 1e9, "unit": "Gcells/s", "cores": int(cores), "kind": "port", "whole_lattice": whole,
            "sample": f"{where} ({cells} cells), {steps} steps after {warmup} warm-up, f64 AoS restatement of FdtdCpuBackend (oracle/), OpenMP on {cores} threads",
            "seconds": secs, "ms_per_step": 1e3 * secs / steps}


def run_reference(args) -> None:
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    case = args.config
    if case != "3":
        w, desc = case_workload(case, args.lattice)
        lattice = w.lattice
        cb = time_cpu_reference(lattice, lattice[2], args.steps, args.warmup, workload=w)
        config = workload_config(args, lattice, 1, desc)
    else:
        lattice = tuple(args.lattice) if args.lattice else ((1024, 1024, 1024) if args.scaling == "strong" else lattice_for(args.gpus))
        cells = lattice[0] * lattice[1] * lattice[2]
        # N=1: the whole lattice, i.e. the same configuration as the GPU arm (18 GB of f64 state at 512^3, ~0.3 s per step on
        # 16 cores); N>1 (up to 1024^3 = 146 GB): a z-slab sample of the N-GPU lattice, named in cpu_baseline.sample.
        whole = args.gpus == 1 and not args.cpu_sample_only and mem_available_bytes() > 1.25 * CPU_BYTES_PER_CELL * cells
        cb = time_cpu_reference(lattice, lattice[2] if whole else args.cpu_planes, args.steps, args.warmup)
        config = workload_config(args, lattice, args.gpus)
    out = {"metric": "Yee cell-updates/sec", "value": cb["value"], "unit": "Gcells/s", "n_gpus": args.gpus, "steps": args.steps,
           "warmup": args.warmup, "ms_per_step": cb["ms_per_step"], "higher_is_better": True, "scaling": args.scaling, "vs_baseline": None,
           "dtype": "f64", "data": "synthetic", "impl": "reference",
           "config": config,
           "cpu_baseline": {k: cb[k] for k in ("value", "unit", "cores", "kind", "sample", "whole_lattice")},
           "arithmetic_note": "the reference's CPU backend computes in f64 (cem-solver/src/fdtd/cpu/mod.rs:299-302); the GPU arm computes in fp32 like the reference's own GPU path (update.wgsl), as north_star asks",
           "e2e": {"value": cb["value"], "unit": "Gcells/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
           "gpu_launches": 0}
    print(json.dumps(out), flush=True)


# ---- multi-GPU parity self-check ------------------------------------------------------------------------------------------

def slab_parity_check(args, rank: int, world: int, local_rank: int, dist, torch) -> dict:
    """SURVEY §8(d) config 4's parity clause, run by the driver's own command: a 256x256x(64·N) heterogeneous lattice with random
    initial fields and a source at a cut, 20 steps in exact arithmetic, decomposed over the N ranks through the SAME transport
    the timed region uses; every rank's slab (SHA-256 of its E and H) against rank 0's single-domain run of the same lattice."""
    import fdtd_b200
    from fdtd_b200.distributed import CudaSlabEngine, P2PSlabStepper, SlabStepper, partition_z

    lattice = (256, 256, 64 * world)
    steps = 20
    parts = partition_z(lattice[2], world)
    zb, ze = parts[rank]
    E, H = fdtd_b200.FieldComponent.E, fdtd_b200.FieldComponent.H

    def planes(z0, z1, comp):
        out = np.empty((z1 - z0, lattice[1], lattice[0], 3), dtype=np.float32)
        for z in range(z0, z1):   # per-plane seeds: every rank can generate exactly its own planes
            out[z - z0] = np.random.default_rng(1000 * comp + z).standard_normal((lattice[1], lattice[0], 3)).astype(np.float32)
        return out

    def workload(z0, z1):
        w = build_workload(lattice, z0, z1)
        cut = parts[0][1]
        w.sources = [((lattice[0] // 2, lattice[1] // 2, cut - 1), w.sources[0][1]), ((3, 5, cut), w.sources[0][1])]
        w.probes = []
        return w

    w = workload(zb, ze)
    tab = w.source_table(steps)
    st = fdtd_b200.FdtdB200Backend(device=local_rank, exact_arithmetic=True, z_slab=(zb, ze), fused_step=None if not args.no_fused else False) \
        .create_instance(w.solver_config(), w.domain()).create_state()
    st.write_field(E, planes(zb, ze, 0), (0, 0, zb), (lattice[0], lattice[1], ze))
    st.write_field(H, planes(zb, ze, 1), (0, 0, zb), (lattice[0], lattice[1], ze))
    st.set_source_table(w.source_points(), tab)
    if args.transport == "p2p" and st.uses_fused and not args.force_nccl:
        stepper = P2PSlabStepper(st, rank, world)
        via = "p2p"
    else:
        stepper = SlabStepper(CudaSlabEngine(st), rank, world, overlap=not args.no_overlap)
        via = "nccl"
    stepper.step(steps)
    st.sync()
    torch.cuda.synchronize()
    mine = [hashlib.sha256(st.read_field(c).tobytes()).hexdigest() for c in (E, H)]
    path = st.step_path
    st.close()
    every = [None] * world
    dist.all_gather_object(every, mine)
    result = {"lattice": list(lattice), "steps": steps, "ranks": world, "transport": via, "arithmetic": "exact", "bit_identical": None}
    if rank == 0:
        wf = workload(0, lattice[2])
        full = fdtd_b200.FdtdB200Backend(device=local_rank, exact_arithmetic=True, fused_step=None if not args.no_fused else False) \
            .create_instance(wf.solver_config(), wf.domain()).create_state()
        full.write_field(E, planes(0, lattice[2], 0))
        full.write_field(H, planes(0, lattice[2], 1))
        full.set_source_table(wf.source_points(), tab)
        full.step(steps)
        ok, nonzero = True, False
        for ci, c in enumerate((E, H)):
            f = full.read_field(c)
            nonzero = nonzero or bool(np.any(f))
            for r, (a, b) in enumerate(parts):
                ok = ok and hashlib.sha256(f[a:b].tobytes()).hexdigest() == every[r][ci]
        full.close()
        result.update({"bit_identical": bool(ok and nonzero), "step_path": path,
                       "compared": "SHA-256 of every rank's E and H slab against the same planes of rank 0's single-domain run"})
    flag = [result.get("bit_identical")]
    dist.broadcast_object_list(flag, src=0)
    result["bit_identical"] = flag[0]
    dist.barrier()
    return result


# ---- the GPU arm -------------------------------------------------------------------------------------------------------------

def run_b200(args) -> None:
    import ctypes as C

    import fdtd_b200
    from fdtd_b200 import _lib as L
    from fdtd_b200.distributed import CudaSlabEngine, SlabStepper, partition_z

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    n_gpus = args.gpus
    if world != n_gpus and world > 1:
        raise SystemExit(f"--gpus {n_gpus} but WORLD_SIZE={world}")
    case = args.config
    if case != "3" and (world > 1 or n_gpus > 1):
        raise SystemExit("--config 1|2|app|thin are single-GPU cases")
    dist = None
    torch = None
    if world > 1:
        import torch
        import torch.distributed as dist

        torch.cuda.set_device(local_rank)
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))

    slab_parity = None
    if world > 1 and not args.no_slab_parity:
        slab_parity = slab_parity_check(args, rank, world, local_rank, dist, torch)
        if slab_parity["bit_identical"] is not True:
            raise SystemExit(f"slab parity self-check FAILED: {slab_parity}")

    description = None
    if case != "3":
        w, description = case_workload(case, args.lattice)
        lattice = w.lattice
        z_begin, z_end = 0, lattice[2]
    else:
        lattice = tuple(args.lattice) if args.lattice else ((1024, 1024, 1024) if args.scaling == "strong" else lattice_for(n_gpus))
        z_begin, z_end = partition_z(lattice[2], world)[rank]
        w = build_workload(lattice, z_begin, z_end)
    nx, ny, nz = lattice
    if args.pml:
        if world > 1 and (args.no_fused or args.transport != "p2p"):
            raise SystemExit("--pml on several GPUs runs inside the fused kernel over the peer-memory transport (drop --no-fused / --transport nccl)")
        w.pml = fdtd_b200.PmlProfiles.graded(w.solver_config(), args.pml, axes=tuple("xyz".index(a) for a in args.pml_axes))
    # the default backend: whatever a caller of the reference-facing API gets (the fused step wherever eligible)
    backend = fdtd_b200.FdtdB200Backend(device=local_rank, exact_arithmetic=args.exact, z_slab=(z_begin, z_end) if world > 1 else None,
                                        fused_step=False if args.no_fused else None)
    inst = backend.create_instance(w.solver_config(), w.domain())
    st = inst.create_state()
    st.register_probes(w.probes)
    stepper = None
    transport = "none"
    if world > 1:
        from fdtd_b200.distributed import P2PSlabStepper

        if args.transport == "p2p" and st.uses_fused:
            # NVLink peer-memory halos (CUDA IPC + device-side flags, stepping inside the C library).  If the box refuses
            # IPC mappings every rank falls back to NCCL together; the line says which transport ran.
            try:
                stepper = P2PSlabStepper(st, rank, world)
                ok = 1.0
            except Exception as exc:  # noqa: BLE001
                print(f"[rank {rank}] p2p transport unavailable: {exc}", file=sys.stderr, flush=True)
                ok = 0.0
            t = torch.tensor([ok], dtype=torch.float64, device="cuda")
            dist.all_reduce(t, op=dist.ReduceOp.MIN)
            if float(t.item()) < 1.0:
                stepper = None
            else:
                transport = "p2p (CUDA IPC peer stores + device flags)"
        if stepper is None:
            if args.pml:
                raise SystemExit("--pml on several GPUs needs the peer-memory transport, which this box refused")
            engine = CudaSlabEngine(st)
            stepper = SlabStepper(engine, rank, world, overlap=not args.no_overlap)
            transport = "nccl send/recv"

    def advance(n):
        if stepper is None:
            st.step(n)
        else:
            stepper.step(n)

    def barrier():
        st.sync()
        if dist is not None:
            torch.cuda.synchronize()
            dist.barrier()

    def max_over_ranks(values):
        if dist is None:
            return list(values)
        t = torch.tensor(list(values), dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return [float(v) for v in t.tolist()]

    K, W = args.steps, args.warmup
    pts = w.source_points()
    cells_total = nx * ny * nz

    # ---- device-resident throughput: forcing table already in HBM, no host traffic in the timed region ----
    st.set_source_table(pts, w.source_table(W))
    advance(W)
    barrier()
    # one untimed probe pass sizes the number of rounds: at least 5 (best of 5), more when K steps are short, so
    # that the timed passes (three kinds per round, below) span >= ~0.6 s and the clock sampler sees them
    st.set_source_table(pts, w.source_table(K, t0=st.time()))
    st.timer_start()
    advance(K)
    probe_ms = max_over_ranks([st.timer_stop()])[0]
    barrier()
    P = args.passes if args.passes > 0 else int(min(max(5, math.ceil(600.0 / max(3.3 * probe_ms, 1e-3))), 60))
    sampler = ClockSampler(local_rank)
    if rank == 0:
        sampler.start()
        time.sleep(0.25)
    cells_arr = np.array([w.cell(p) for p in pts], dtype=np.uint64)
    n_src, n_pr = len(pts), len(w.probes)
    probe_out = np.zeros((1, max(n_pr, 1), 6), dtype=np.float32)
    probe_ptr = probe_out.ctypes.data_as(C.POINTER(C.c_float))
    lib = fdtd_b200.lib()

    # the K steps' SourceData records of an end-to-end pass are laid out before its clock starts (evaluating the source
    # function is host work of the caller, source.rs:39-98); what is timed per step is the hand-over (H2D), the step and the
    # probe read (D2H)
    def source_records():
        tab32 = w.source_table(K, t0=st.time()).astype(np.float32)
        out = []
        for s in range(K):
            arr = (L.Source * n_src)()
            for k in range(n_src):
                arr[k].cell = int(cells_arr[k])
                arr[k].j = (C.c_float * 3)(*tab32[s, k, :3])
                arr[k].m = (C.c_float * 3)(*tab32[s, k, 3:])
            out.append(arr)
        return out

    gap_s = max(args.pass_gap_ms, 0) * 1e-3

    def device_pass():
        """K steps on device-resident forcing, one CUDA-event pair around them → ms."""
        st.set_source_table(pts, w.source_table(K, t0=st.time()))
        barrier()
        if gap_s:
            time.sleep(gap_s)   # idle gap: every pass starts from the power state the peak (a best-of-10 burst copy) was measured in
        st.timer_start()
        advance(K)
        ms_p = st.timer_stop()
        barrier()
        return ms_p

    def e2e_pass(lagging: bool):
        """K x [set_sources (H2D) → step(1) → read_probes (D2H)] through the reference-facing C ABI with host buffers → seconds.
        lagging: the row of step s is read (through the event behind its copy, fdtd_b200_read_probes) while step s + 1 runs, so
        the stream never drains; reported beside the strict number, not as it."""
        steps_src = source_records()
        tick0 = st.tick()
        barrier()
        if gap_s:
            time.sleep(gap_s)
        c0 = time.perf_counter()
        for s in range(K):
            L.check(lib.fdtd_b200_set_sources(st._h, n_src, steps_src[s]), st._h)
            advance(1)
            if not lagging:
                L.check(lib.fdtd_b200_read_probes(st._h, tick0 + s, 1, probe_ptr), st._h)
            elif s:
                L.check(lib.fdtd_b200_read_probes(st._h, tick0 + s - 1, 1, probe_ptr), st._h)
        if lagging:
            L.check(lib.fdtd_b200_read_probes(st._h, tick0 + K - 1, 1, probe_ptr), st._h)
        barrier()
        return time.perf_counter() - c0

    # The three kinds of pass alternate in EVERY round — device-resident, end to end, end to end with the lagging read — so that
    # clock drift under the power cap (a B200 running this HBM-bound kernel reaches its cap after 0.3 … 1 s, box by box) hits all
    # of them alike, and each reports the median of its P passes.  (Until round 2 the end-to-end passes ran in the first 9 rounds
    # only: on a box that throttled from round 12 on, `e2e` came out above `value`.)  Three of the rounds add a pass with an event pair around every kernel launch:
    # the dominant kernel's average duration for the roofline.  That instrumentation is kept out of the `value` passes: on
    # several GPUs the events sit between the boundary- and interior-stream launches of consecutive steps and cost 3 %
    # (N=2: 249 against 256 Gcells/s; N=1: 0.2 %).
    EP = P
    instrumented_rounds = {0, EP // 2, EP - 1}
    t0 = time.monotonic()
    pass_ms, e2e_passes, pipe_passes, prof_runs = [], [], [], []
    launches = 0
    for p in range(P):
        launches0 = st.kernel_launches
        pass_ms.append(device_pass())
        launches = st.kernel_launches - launches0
        if p < EP:
            e2e_passes.append(e2e_pass(False))
            pipe_passes.append(e2e_pass(True))
            if p in instrumented_rounds:
                st.set_profiling(True)
                ms_i = device_pass()
                prof_runs.append((ms_i, st.profile(0), st.profile(1)))
                st.set_profiling(False)
    pass_ms = max_over_ranks(pass_ms)
    ms = statistics.median(pass_ms)
    value = cells_total * K / (ms * 1e-3) / 1e9
    prof_runs.sort(key=lambda r: r[0])
    ms_instrumented, prof0, prof1 = prof_runs[len(prof_runs) // 2]      # the median instrumented pass
    prof = [prof0, prof1]
    t1 = time.monotonic()   # clocks are summarised over every timed region (device-resident passes + end-to-end passes)
    if rank == 0:
        sampler.stop()
    e2e_s = statistics.median(max_over_ranks(e2e_passes))
    pipe_s = statistics.median(max_over_ranks(pipe_passes))
    e2e_value = cells_total * K / e2e_s / 1e9

    if rank != 0:
        if dist is not None:
            dist.destroy_process_group()
        return

    peak, peak_src = read_peak_hbm()
    # dominant kernel: the fused single-pass step kernel (6 fp32 reads + 6 fp32 writes + 1 B index = 49 B per cell it
    # updates), or — on the two-kernel path — the slower of the two half-step kernels (37 B per cell each)
    fused = st.uses_fused
    kern = []
    names = ((0, "step_fused", BYTES_PER_CELL_FUSED),) if fused else ((0, "update_h", BYTES_PER_CELL_HALF_STEP), (1, "update_e", BYTES_PER_CELL_HALF_STEP))
    # with --pml every ψ value of the slabs (4 per slab cell and axis) is read and written once per step on top of that
    psi_bytes = 8 * int(lib.fdtd_b200_pml_state_floats(st._h)) if args.pml else 0
    for which, name, bpc in names:
        tot_ms, n_l, n_cells = prof[which]
        if n_l:
            extra = psi_bytes * n_l / (1 if fused else 2)
            kern.append({"kernel": name, "avg_ms": tot_ms / n_l, "launches": int(n_l), "algorithmic_bytes_per_cell": bpc,
                         "achieved": (bpc * n_cells + extra) / (tot_ms * 1e-3) / 1e9})
    dom = min(kern, key=lambda k: k["achieved"])
    step_bytes_per_gpu = (BYTES_PER_CELL_FUSED if fused else BYTES_PER_CELL_STEP) * cells_total / n_gpus + psi_bytes
    if world > 1:
        # a slab step is two or three launches on two streams that OVERLAP (boundary planes first, interior concurrently):
        # summing their durations double-counts, so the N>1 roofline is the whole step's bytes over the whole step's device time
        achieved = step_bytes_per_gpu / (ms / K * 1e-3) / 1e9
        how = f"per GPU: algorithmic bytes of one step / median device time of one step ({ms / K:.4f} ms; boundary and interior launches overlap, their per-launch times are listed under kernels but not summed)"
        avg_ms = ms / K
        alg_launch = step_bytes_per_gpu
    else:
        achieved = dom["achieved"]
        how = (f"the median of {len(prof_runs)} instrumented passes of {K} steps (an event pair around every launch; {ms_instrumented / K:.4f} ms per step), "
               "interleaved with the timed passes")
        avg_ms = dom["avg_ms"]
        alg_launch = dom["algorithmic_bytes_per_cell"] * cells_total + psi_bytes / (1 if fused else 2)
    traffic = None
    try:  # per-launch DRAM bytes of the dominant kernel from the committed ncu capture (same lattice only)
        with open(os.path.join(ROOT, "profiles", "ncu_traffic.json")) as f:
            tr = json.load(f)
        other = tr.get("other_lattices", {}).get("x".join(str(v) for v in lattice))
        if n_gpus == 1 and other is not None and not args.pml and other["kernel"] in st.step_path:
            traffic = other["dram_read_bytes"] + other["dram_write_bytes"]
        elif n_gpus == 1 and list(lattice) == tr["lattice"]:
            if fused:
                key = "k_step_fused" + (f"_pml{args.pml}" if args.pml else "")
            else:
                key = ("k_update_h" if dom["kernel"] == "update_h" else "k_update_e") + ("_bulk" if st.uses_bulk else "_vec4")
            traffic = tr[key]["dram_read_bytes"] + tr[key]["dram_write_bytes"]
    except Exception:
        traffic = None
    roofline = {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                "traffic": traffic, "traffic_unit": "bytes per launch (ncu dram__bytes_read.sum + dram__bytes_write.sum; committed capture profiles/ncu_traffic.json, not this run)",
                "algorithmic_bytes_per_launch": alg_launch, "kernel": (dom["kernel"] + (" (k_step_2d)" if "k_step_2d" in st.step_path else "")) if world == 1 else "step (all launches of one step)",
                "kernel_avg_from": how, "kernel_avg_ms": avg_ms, "peak_source": peak_src,
                "algorithmic_bytes_per_cell": dom["algorithmic_bytes_per_cell"] if world == 1 else (BYTES_PER_CELL_FUSED if fused else BYTES_PER_CELL_STEP),
                # the same run against SURVEY §8d's two-pass figure (74 B per cell update): a fused pass can exceed 1.0 here
                "step_equivalent_74B_gbs": BYTES_PER_CELL_STEP * value / n_gpus, "step_equivalent_74B_frac": BYTES_PER_CELL_STEP * value / n_gpus / peak,
                "kernels": kern}
    if 25 * cells_total / n_gpus < 126e6:
        roofline["note"] = "the lattice state fits the 126 MB L2: this case is bound by launch / grid-synchronisation latency, not by HBM; frac is reported for completeness"
    cpu_baseline = None
    if n_gpus == 1 and not args.no_cpu_baseline:
        if case != "3":
            # the small cases run the CPU leg on the SAME lattice, bounded to ~10-20 s of CPU work
            cpu_steps = int(min(max(2e9 / max(cells_total, 1), 4), 2000)) if args.cpu_steps <= 0 else args.cpu_steps
            cb = time_cpu_reference(lattice, nz, cpu_steps, 2, workload=w)
            cpu_baseline = {k: cb[k] for k in ("value", "unit", "cores", "kind", "sample", "whole_lattice")}
            c1 = time_cpu_reference(lattice, nz, max(cpu_steps // 8, 2), 1, threads=1, workload=w)
            cpu_baseline["single_thread"] = {k: c1[k] for k in ("value", "unit", "cores", "sample")}
        else:
            cpu_steps = args.cpu_steps if args.cpu_steps > 0 else 10
            cb = time_cpu_reference(lattice, args.cpu_planes, cpu_steps, 2)
            cpu_baseline = {k: cb[k] for k in ("value", "unit", "cores", "kind", "sample", "whole_lattice")}
            # SURVEY §8d's second row: one thread ≙ FdtdCpuBackend::single_threaded (cpu/mod.rs:187-193), on a smaller sample
            c1 = time_cpu_reference(lattice, max(args.cpu_planes // 4, 2), max(cpu_steps // 2, 2), 1, threads=1)
            cpu_baseline["single_thread"] = {k: c1[k] for k in ("value", "unit", "cores", "sample")}
    out = {"metric": "Yee cell-updates/sec", "value": value, "unit": "Gcells/s", "n_gpus": n_gpus, "steps": K, "warmup": W,
           "ms_per_step": ms / K, "higher_is_better": True, "scaling": args.scaling, "vs_baseline": None, "dtype": "f32", "data": "synthetic",
           "config": workload_config(args, lattice, n_gpus, description),
           "timing": {"passes": P, "steps_per_pass": K, "statistic": "median over passes of the per-pass max over ranks (CUDA events on the launching stream)",
                      "pass_ms": [round(v, 5) for v in pass_ms], "idle_gap_before_each_pass_ms": args.pass_gap_ms, "best_value": cells_total * K / (min(pass_ms) * 1e-3) / 1e9,
                      "worst_value": cells_total * K / (max(pass_ms) * 1e-3) / 1e9},
           "impl_detail": {"row_pitch": st.row_pitch, "arithmetic": "exact" if args.exact else "fast", "step_path": st.step_path,
                           "step_kernel": "fused single pass (ping-pong buffers)" if st.uses_fused else "two half-step kernels (in place)",
                           "backend": "FdtdB200Backend() defaults" + (" + in-place step (--no-fused)" if args.no_fused else ""),
                           "device_bytes_in_use": st.memory_in_use,
                           "halo_overlap": bool(world > 1 and not args.no_overlap), "halo_transport": transport,
                           **({"cpml_cells_per_face": args.pml, "cpml": "inside the fused step kernel" if st.pml_in_kernel() else "psi passes over the slabs after the step kernel(s); kernel times include them",
                               "cpml_psi_bytes_per_step": psi_bytes} if args.pml else {})},
           "roofline": roofline, "cpu_baseline": cpu_baseline,
           "e2e": {"value": e2e_value, "unit": "Gcells/s", "h2d_bytes_per_step": 32 * n_src, "d2h_bytes_per_step": 24 * n_pr,
                   "ms_per_step": 1e3 * e2e_s / K, "passes": EP, "statistic": "median pass (host clock around K x [set_sources -> step(1) -> read_probes], max over ranks)",
                   "read_lagging_one_step": {"value": cells_total * K / pipe_s / 1e9, "ms_per_step": 1e3 * pipe_s / K,
                                             "note": "same per-step H2D forcing and D2H probe row; the row of step s is read while step s+1 runs"}},
           "gpu_launches": int(launches), "clocks": sampler.summary(t0, t1)}
    if slab_parity is not None:
        out["slab_parity"] = slab_parity
    print(json.dumps(out), flush=True)
    if dist is not None:
        dist.destroy_process_group()


def main() -> None:
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=200)
    ap.add_argument("--warmup", type=int, default=10)
    ap.add_argument("--passes", type=int, default=0, help="timed passes of --steps steps each (0 = auto: >= 5, more for short passes)")
    ap.add_argument("--pass-gap-ms", type=float, default=0.0, help="idle time before every timed pass (outside the timed region)")
    ap.add_argument("--impl", choices=("b200", "reference"), default="b200")
    ap.add_argument("--config", choices=("1", "2", "3", "app", "thin"), default="3",
                    help="3 = BASELINE configs[2], the headline (default); 1 = 64^3 vacuum; 2 = 256^3 inclusions; app = 100x100x1; thin = 4096x4096x1")
    ap.add_argument("--lattice", type=lambda s: [int(v) for v in s.split(",")], default=None)
    ap.add_argument("--exact", action="store_true", help="FDTD_B200_EXACT_ARITHMETIC kernels")
    ap.add_argument("--no-overlap", action="store_true")
    ap.add_argument("--transport", choices=("p2p", "nccl"), default="p2p", help="halo transport for N>1 (p2p needs the fused step)")
    ap.add_argument("--force-nccl", action="store_true", help=argparse.SUPPRESS)
    ap.add_argument("--no-fused", action="store_true", help="two-kernel path (in-place, 24 B/cell of field memory) instead of the fused step")
    ap.add_argument("--scaling", choices=("weak", "strong"), default="weak",
                    help="weak: 512^3 cells per GPU (default, the driver's scaling run); strong: one fixed lattice (1024^3 unless --lattice) for every N")
    ap.add_argument("--pml-axes", default="xyz", help="axes that carry the --pml slabs (experiments: x, y, z, xy, ...)")
    ap.add_argument("--pml", type=int, default=0, help="add CPML slabs of this many cells on all six faces (not the headline workload)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-slab-parity", action="store_true", help="skip the N>1 slab-parity self-check before the timed region")
    ap.add_argument("--cpu-planes", type=int, default=32, help="z-planes in the CPU legs' sample slab")
    ap.add_argument("--cpu-steps", type=int, default=0, help="steps of the cpu_baseline leg (0 = auto)")
    ap.add_argument("--cpu-sample-only", action="store_true", help="--impl reference at N=1: time the z-slab sample instead of the whole lattice")
    args = ap.parse_args()
    if args.warmup < 3:
        args.warmup = 3
    if args.transport == "nccl":
        args.force_nccl = True
    if args.impl == "reference":
        run_reference(args)
    else:
        run_b200(args)


if __name__ == "__main__":
    main()
