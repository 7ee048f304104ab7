#!/usr/bin/env python
"""Benchmark of the FDTD time-stepping path (BASELINE.json metric: Yee cell-updates/s in Gcells/s).

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl b200|reference] [--lattice nx,ny,nz]

A "step" is one full leapfrog step (H half-step then E half-step, six components) of the whole lattice.
N = 1 : BASELINE configs[2] — the 512^3 heterogeneous lattice inside the lossy absorbing layer (SURVEY §8d
        config 3, the headline single-GPU roofline run).
N > 1 : the same kind of lattice grown so every GPU keeps 512^3 cells (weak scaling), z-slab decomposed:
        N=2 1024x512x512, N=4 1024x1024x512, N=8 1024^3; halo planes stored into the neighbours' ghost planes over NVLink peer
        memory with device-side flags (--transport p2p, default) or exchanged with NCCL send/recv (--transport nccl).
`--impl reference` times the CPU restatement of the reference's own CPU backend (oracle/, f64 AoS, all host
threads) on a bounded z-slab sample of the same lattice; the Rust reference itself cannot be built here.
Prints ONE JSON line on rank 0.
"""
from __future__ import annotations

import argparse
import json
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


def lattice_for(n_gpus: int) -> tuple:
    return {1: (512, 512, 512), 2: (1024, 512, 512), 4: (1024, 1024, 512), 8: (1024, 1024, 1024)}.get(
        n_gpus, (512, 512, 512 * n_gpus))


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
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.FIELDS}", "--format=csv,noheader,nounits", "-lms", "50", "-i", str(self.gpu)],
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


def build_workload(lattice, z_begin, z_end):
    from fdtd_b200 import configs
    from fdtd_b200.solver import GaussianPulse

    pal = configs.heterogeneous_palette()
    ids = configs.heterogeneous_slab_labels(lattice, z_begin, z_end)
    nx, ny, nz = lattice
    src = GaussianPulse(12.5, 2.5).with_amplitudes((0.0, 0.0, 50.0), (0.0, 0.0, 0.0))
    sources = [((nx // 2, ny // 2, nz // 2), src)]
    probes = [(min(nx // 2 + 16 * (k + 1), nx - 1), ny // 2, nz // 2) for k in range(4)] + [(nx // 2, min(ny // 2 + 16 * (k + 1), ny - 1), nz // 2) for k in range(4)]
    w = configs.Workload(f"heterogeneous{nx}x{ny}x{nz}", tuple(lattice), configs.REDUCED_RES, configs.PhysicalConstants.REDUCED,
                         pal, ids, sources, probes, 0)
    return w


def cpu_sample(lattice, planes: int):
    """A z-slab sample of the benchmark lattice for the CPU legs: `planes` planes around the source plane."""
    nx, ny, nz = lattice
    z0 = max(nz // 2 - planes // 2, 0)
    z1 = min(z0 + planes, nz)
    return z0, z1


def time_cpu_reference(lattice, planes: int, steps: int, warmup: int, threads=None) -> dict:
    """Times the oracle's f64 restatement of FdtdCpuBackend (multi-threaded ≙ cpu/mod.rs:148-157) on a slab sample."""
    import oracle
    from fdtd_b200 import configs

    oracle.build()
    z0, z1 = cpu_sample(lattice, planes)
    nx, ny, nz = lattice
    sub = (nx, ny, z1 - z0)
    ids = configs.heterogeneous_slab_labels(lattice, z0, z1)
    pal = np.array([m.as_row() for m in configs.heterogeneous_palette()], dtype=np.float64)
    cores = oracle.max_threads(fast=True) if threads is None else threads
    oracle.set_threads(cores, fast=True)
    s = oracle.RefSolver(sub, (1.0, 1.0, 1.0), 0.25, precision=oracle.F64, fast=True)
    s.set_materials(pal, ids, 1.0, 1.0)
    w = build_workload(lattice, z0, z1)
    pts = [(nx // 2, ny // 2, nz // 2 - z0)]
    tab = w.source_table(warmup + steps)
    s.run(warmup, pts, tab[:warmup])
    _, secs = s.run(steps, pts, tab[warmup:])
    cells = nx * ny * (z1 - z0)
    return {"value": cells * steps / secs / 1e9, "unit": "Gcells/s", "cores": int(cores), "kind": "port",
            "sample": f"planes [{z0},{z1}) of the {nx}x{ny}x{nz} lattice ({cells} cells), {steps} steps, f64 AoS restatement of FdtdCpuBackend (oracle/), OpenMP",
            "seconds": secs, "ms_per_step": 1e3 * secs / steps}


def run_reference(args) -> None:
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    lattice = tuple(args.lattice) if args.lattice else lattice_for(args.gpus)
    cb = time_cpu_reference(lattice, args.cpu_planes, args.steps, args.warmup)
    out = {"metric": "Yee cell-updates/sec", "value": cb["value"], "unit": "Gcells/s", "n_gpus": args.gpus, "steps": args.steps,
           "warmup": args.warmup, "ms_per_step": cb["ms_per_step"], "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
           "dtype": "f64", "data": "synthetic", "impl": "reference",
           "config": {"workload": f"heterogeneous {lattice[0]}x{lattice[1]}x{lattice[2]} lattice + 16-cell lossy layer (BASELINE configs[2] shape), CPU leg on a bounded z-slab sample",
                      "lattice": list(lattice)},
           "cpu_baseline": {k: cb[k] for k in ("value", "unit", "cores", "kind", "sample")},
           "e2e": {"value": cb["value"], "unit": "Gcells/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
           "gpu_launches": 0}
    print(json.dumps(out), flush=True)


def run_b200(args) -> None:
    import fdtd_b200
    from fdtd_b200 import _lib as L
    from fdtd_b200.distributed import CudaSlabEngine, SlabStepper, partition_z

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    n_gpus = args.gpus
    if world != n_gpus and world > 1:
        raise SystemExit(f"--gpus {n_gpus} but WORLD_SIZE={world}")
    dist = None
    torch = None
    if world > 1:
        import torch
        import torch.distributed as dist

        torch.cuda.set_device(local_rank)
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))

    lattice = tuple(args.lattice) if args.lattice else ((1024, 1024, 1024) if args.scaling == "strong" else lattice_for(n_gpus))
    nx, ny, nz = lattice
    z_begin, z_end = partition_z(nz, world)[rank]
    w = build_workload(lattice, z_begin, z_end)
    if args.pml:
        if world > 1 and (args.no_fused or args.transport != "p2p"):
            raise SystemExit("--pml on several GPUs runs inside the fused kernel over the peer-memory transport (drop --no-fused / --transport nccl)")
        w.pml = fdtd_b200.PmlProfiles.graded(w.solver_config(), args.pml)
    backend = fdtd_b200.FdtdB200Backend(device=local_rank, exact_arithmetic=args.exact, z_slab=(z_begin, z_end) if world > 1 else None,
                                        fused_step=not args.no_fused)
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

    K, W = args.steps, args.warmup
    table = w.source_table(W + 4 * K + 8)
    pts = w.source_points()

    # ---- device-resident throughput: forcing table already in HBM, no host traffic in the timed region ----
    st.set_source_table(pts, table)
    advance(W)
    barrier()
    sampler = ClockSampler(local_rank)
    if rank == 0:
        sampler.start()
        time.sleep(0.3)
    # Two passes of K steps.  The first is timed as a whole (one event pair around fdtd_b200_step(h, K)) and gives `value`; the
    # second carries an event pair around every kernel launch and gives the dominant kernel's average duration for the
    # roofline.  The per-launch events are kept out of the first pass because on several GPUs they sit between the boundary-
    # and interior-stream launches of consecutive steps and cost 3 % (N=2: 249 against 256 Gcells/s; N=1: 0.2 %).
    launches0 = st.kernel_launches
    barrier()
    t0 = time.monotonic()
    st.timer_start()
    advance(K)
    ms = st.timer_stop()
    barrier()
    launches = st.kernel_launches - launches0
    st.set_profiling(True)
    barrier()
    st.timer_start()
    advance(K)
    ms_instrumented = st.timer_stop()
    barrier()
    prof = [st.profile(0), st.profile(1)]
    st.set_profiling(False)
    if dist is not None:
        t = torch.tensor([ms], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms = float(t.item())
    cells_total = nx * ny * nz
    value = cells_total * K / (ms * 1e-3) / 1e9

    # ---- end to end through the reference-facing update pass: per step the host hands over that step's forcing
    #      (H2D), the engine steps once, and the step's probe samples come back to the host (D2H) ----
    st.sync()
    tick0 = st.tick()
    tab32 = table.astype(np.float32)
    cells_arr = np.array([w.cell(p) for p in pts], dtype=np.uint64)
    n_src, n_pr = len(pts), len(w.probes)
    probe_out = np.zeros((1, n_pr, 6), dtype=np.float32)
    import ctypes as C

    lib = fdtd_b200.lib()
    # the K steps' SourceData records are laid out before the clock starts (evaluating the source function is host work of
    # the caller, source.rs:39-98); what is timed per step is the hand-over (H2D), the step and the probe read (D2H)
    def source_records(first_tick):
        out = []
        for s in range(K):
            row = tab32[first_tick + s]
            arr = (L.Source * n_src)()
            for k in range(n_src):
                arr[k].cell = int(cells_arr[k])
                arr[k].j = (C.c_float * 3)(*row[k, :3])
                arr[k].m = (C.c_float * 3)(*row[k, 3:])
            out.append(arr)
        return out

    steps_src = source_records(tick0)
    probe_ptr = probe_out.ctypes.data_as(C.POINTER(C.c_float))
    barrier()
    e0 = time.perf_counter()
    for s in range(K):
        L.check(lib.fdtd_b200_set_sources(st._h, n_src, steps_src[s]), st._h)
        advance(1)
        L.check(lib.fdtd_b200_read_probes(st._h, tick0 + s, 1, probe_ptr), st._h)
    barrier()
    e2e_s = time.perf_counter() - e0
    # the same protocol with the probe read lagging one step: the row of step s is read (through the event behind its copy,
    # fdtd_b200_read_probes) while step s + 1 runs, so the stream never drains.  Reported beside the strict number, not as it.
    tick1 = st.tick()
    steps_src = source_records(tick1)
    barrier()
    p0 = time.perf_counter()
    for s in range(K):
        L.check(lib.fdtd_b200_set_sources(st._h, n_src, steps_src[s]), st._h)
        advance(1)
        if s:
            L.check(lib.fdtd_b200_read_probes(st._h, tick1 + s - 1, 1, probe_ptr), st._h)
    L.check(lib.fdtd_b200_read_probes(st._h, tick1 + K - 1, 1, probe_ptr), st._h)
    barrier()
    pipe_s = time.perf_counter() - p0
    t1 = time.monotonic()   # clocks are summarised over both timed regions (device-resident + end-to-end)
    if rank == 0:
        sampler.stop()
    if dist is not None:
        t = torch.tensor([e2e_s], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        e2e_s = float(t.item())
        t = torch.tensor([pipe_s], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        pipe_s = float(t.item())
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
    traffic = None
    try:  # per-launch DRAM bytes of the dominant kernel from the committed ncu capture (same lattice only)
        with open(os.path.join(ROOT, "profiles", "ncu_traffic.json")) as f:
            tr = json.load(f)
        if n_gpus == 1 and list(lattice) == tr["lattice"]:
            if fused:
                key = "k_step_fused" + (f"_pml{args.pml}" if args.pml else "")
            else:
                key = ("k_update_h" if dom["kernel"] == "update_h" else "k_update_e") + ("_bulk" if st.uses_bulk else "_vec4")
            traffic = tr[key]["dram_read_bytes"] + tr[key]["dram_write_bytes"]
    except Exception:
        traffic = None
    roofline = {"bound": "hbm", "achieved": dom["achieved"], "peak": peak, "unit": "GB/s", "frac": dom["achieved"] / peak,
                "traffic": traffic, "traffic_unit": "bytes per launch (ncu dram__bytes_read.sum + dram__bytes_write.sum)",
                "algorithmic_bytes_per_launch": dom["algorithmic_bytes_per_cell"] * cells_total / n_gpus + psi_bytes / (1 if fused else 2), "kernel": dom["kernel"], "kernel_avg_from": f"a second pass of {K} steps with an event pair around every launch ({ms_instrumented / K:.4f} ms per step)",
                "kernel_avg_ms": dom["avg_ms"], "peak_source": peak_src, "algorithmic_bytes_per_cell": dom["algorithmic_bytes_per_cell"],
                # the same run against SURVEY §8d's two-pass figure (74 B per cell update): a fused pass can exceed 1.0 here
                "step_equivalent_74B_gbs": BYTES_PER_CELL_STEP * value / n_gpus, "step_equivalent_74B_frac": BYTES_PER_CELL_STEP * value / n_gpus / peak,
                "kernels": kern}
    cpu_baseline = None
    if n_gpus == 1 and not args.no_cpu_baseline:
        cb = time_cpu_reference(lattice, args.cpu_planes, args.cpu_steps, 2)
        cpu_baseline = {k: cb[k] for k in ("value", "unit", "cores", "kind", "sample")}
        # SURVEY §8d's second row: one thread ≙ FdtdCpuBackend::single_threaded (cpu/mod.rs:187-193), on a smaller sample
        c1 = time_cpu_reference(lattice, max(args.cpu_planes // 4, 2), max(args.cpu_steps // 2, 2), 1, threads=1)
        cpu_baseline["single_thread"] = {k: c1[k] for k in ("value", "unit", "cores", "sample")}
    out = {"metric": "Yee cell-updates/sec", "value": value, "unit": "Gcells/s", "n_gpus": n_gpus, "steps": K, "warmup": W,
           "ms_per_step": ms / K, "higher_is_better": True, "scaling": args.scaling, "vs_baseline": None, "dtype": "f32", "data": "synthetic",
           "config": {"workload": f"heterogeneous {nx}x{ny}x{nz} lattice (64 balls of 4 material kinds) inside the 16-cell graded lossy layer, "
                                  "Gaussian point source, 8 probes (BASELINE configs[2]; " + ("the same lattice for every N: strong scaling, BASELINE configs[3])" if args.scaling == "strong" else "weak-scaled 512^3 cells per GPU for N>1)"),
                      "lattice": [nx, ny, nz], "cells_per_gpu": cells_total // n_gpus, "decomposition": f"z-slabs x{n_gpus}",
                      "l2": f"working set {49 if st.uses_fused else 25} B/cell x {cells_total // n_gpus / 1e6:.0f}M cells per GPU >> 126 MB L2 (no flush needed)",
                      "row_pitch": st.row_pitch, "arithmetic": "exact" if args.exact else "fast", "step_kernel": "fused single pass (ping-pong buffers)" if st.uses_fused else "two half-step kernels (in place)", "halo_overlap": bool(world > 1 and not args.no_overlap), "halo_transport": transport,
                      **({"cpml_cells_per_face": args.pml, "cpml": "inside the fused step kernel" if st.pml_in_kernel() else "psi passes over the slabs after the step kernel(s); kernel times include them",
                         "cpml_psi_bytes_per_step": psi_bytes} if args.pml else {})},
           "roofline": roofline, "cpu_baseline": cpu_baseline,
           "e2e": {"value": e2e_value, "unit": "Gcells/s", "h2d_bytes_per_step": 32 * n_src, "d2h_bytes_per_step": 24 * n_pr,
                   "ms_per_step": 1e3 * e2e_s / K,
                   "read_lagging_one_step": {"value": cells_total * K / pipe_s / 1e9, "ms_per_step": 1e3 * pipe_s / K,
                                             "note": "same per-step H2D forcing and D2H probe row; the row of step s is read while step s+1 runs"}},
           "gpu_launches": int(launches), "clocks": sampler.summary(t0, t1)}
    print(json.dumps(out), flush=True)
    if dist is not None:
        dist.destroy_process_group()


def main() -> None:
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=200)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", choices=("b200", "reference"), default="b200")
    ap.add_argument("--lattice", type=lambda s: [int(v) for v in s.split(",")], default=None)
    ap.add_argument("--exact", action="store_true", help="FDTD_B200_EXACT_ARITHMETIC kernels")
    ap.add_argument("--no-overlap", action="store_true")
    ap.add_argument("--transport", choices=("p2p", "nccl"), default="p2p", help="halo transport for N>1 (p2p needs the fused step)")
    ap.add_argument("--no-fused", action="store_true", help="two-kernel path (in-place, 24 B/cell of field memory) instead of the fused step")
    ap.add_argument("--scaling", choices=("weak", "strong"), default="weak",
                    help="weak: 512^3 cells per GPU (default, the driver's scaling run); strong: one fixed lattice (1024^3 unless --lattice) for every N")
    ap.add_argument("--pml", type=int, default=0, help="add CPML slabs of this many cells on all six faces (not the headline workload)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--cpu-planes", type=int, default=32, help="z-planes in the CPU legs' sample slab")
    ap.add_argument("--cpu-steps", type=int, default=10)
    args = ap.parse_args()
    if args.warmup < 3:
        args.warmup = 3
    if args.impl == "reference":
        run_reference(args)
    else:
        run_b200(args)


if __name__ == "__main__":
    main()
