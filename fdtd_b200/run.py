"""Config-file driven runs: `python -m fdtd_b200.run config.json [--steps N] [--batched] [--out DIR]`.

`prepare` is `RunFdtd::run_fdtd_with_backend` (cem-app/src/solver/runner.rs:247-380) without the window: volume → AABB →
`FdtdSolverConfig` → lattice, Courant warning, `memory_required` against `memory_limit` (bail like runner.rs:304-330),
`CoordinateTransformations::for_fdtd`, the scene's colliders rasterised into the material lattice on the device, sources
snapped to their nearest lattice point (runner.rs:852-869), one projection per observer (runner.rs:732-818; the target is a
`lattice.xy` RGBA8 image — in memory, a PNG sequence or a GIF instead of a render texture).  `run` is the solver thread's
loop (runner.rs:471-557): stop condition, update pass, observers on the WALL-CLOCK cadence `observation_delay`
(runner.rs:464, 530-542: 40 ms = 25 pictures per second, whatever the step rate), optional `step_delay`.  `--batched` hands the
steps between two observations to the device in one call (`fdtd_b200_step(h, n)`, device-resident forcing table; on lattices
that stay on chip that is ONE kernel launch for the n steps) and observes every `--observe-every` steps instead.
"""
from __future__ import annotations

import argparse
import json
import math
import sys
import time as _time
import warnings
from dataclasses import dataclass, field
from typing import Callable, Optional

import numpy as np

from . import _lib as L
from . import config as cfgmod
from . import io as fio
from .driver import Sources, StopCondition
from .scene import CoordinateTransformations, PmlRegion, SceneDomain, Shape, quaternion_to_matrix
from .solver import (B200Projection, FdtdB200Backend, FdtdSolverConfig, FieldComponent, ImageTarget, ProjectionParameters)

APP_COLOR_MAP_CODE = "var x = value.z"   # marker of the application's colour-map snippet (runner.rs:761-780)


@dataclass
class Observer:
    spec: cfgmod.ObserverSpec
    projection: B200Projection
    frames: int = 0


@dataclass
class PreparedRun:
    config: cfgmod.SolverConfig
    solver_config: FdtdSolverConfig
    lattice: tuple
    transformations: CoordinateTransformations
    instance: object
    state: object
    sources: Sources
    observers: list
    memory_required: int
    skipped_sources: int = 0
    courant_ok: bool = True
    timings: dict = field(default_factory=dict)

    def close(self) -> None:
        for o in self.observers:
            t = o.projection.target
            if hasattr(t, "close"):
                t.close()
        self.state.close()


def volume_aabb(volume: cfgmod.Volume, entities: list) -> tuple:
    """`Volume::aabb` (config.rs:110-143) → (mins, maxs) as f32.  Fixed: centre ± half extents.  SceneAabb: the box around every
    collider of the scene plus the margin (rotated observers frames are not restated: identity rotation only)."""
    if volume.kind == "Fixed":
        c = np.asarray(volume.isometry.translation, dtype=np.float32)
        h = np.asarray(volume.half_extents, dtype=np.float32)
        return c - h, c + h
    if tuple(volume.rotation()) != (1.0, 0.0, 0.0, 0.0):
        raise cfgmod.ConfigError("Volume::SceneAabb with a rotation is not supported by the headless runner (use Volume::Fixed)")
    lo, hi = None, None
    for e in entities:
        if e.collider is None:
            continue
        kind = e.collider[0]
        half = (np.full(3, e.collider[1]) if kind == "Ball" else np.asarray(e.collider[1]) if kind == "Cuboid"
                else np.array([e.collider[2], e.collider[1], e.collider[2]]))
        ext = np.abs(quaternion_to_matrix(e.transform.rotation)) @ half.astype(np.float64)
        c = np.asarray(e.transform.translation, dtype=np.float64)
        lo = c - ext if lo is None else np.minimum(lo, c - ext)
        hi = c + ext if hi is None else np.maximum(hi, c + ext)
    if lo is None:
        raise cfgmod.ConfigError("invalid aabb: Volume::SceneAabb of a scene without colliders (runner.rs:273-275 bails)")
    m = np.asarray(volume.margin, dtype=np.float64)
    return (lo - m).astype(np.float32), (hi + m).astype(np.float32)


def _shape(e: cfgmod.SceneEntity) -> Shape:
    kind = e.collider[0]
    mat = e.material
    if kind == "Ball":
        return Shape.ball(e.collider[1], mat, e.transform.translation)
    if kind == "Cuboid":
        return Shape.cuboid(e.collider[1], mat, e.transform.translation, e.transform.rotation)
    return Shape.cylinder(e.collider[1], e.collider[2], mat, e.transform.translation, e.transform.rotation)


def prepare(cfg: cfgmod.SolverConfig, out_dir: Optional[str] = None, exact_arithmetic: bool = False, device: Optional[int] = None) -> PreparedRun:
    t0 = _time.monotonic()
    common, fdtd = cfg.common, cfg.fdtd
    mins, maxs = volume_aabb(common.volume, cfg.scene)
    size = (maxs - mins).astype(np.float64)
    if not np.all(np.isfinite(size)) or np.any(size < 0):
        raise cfgmod.ConfigError(f"invalid aabb: mins {mins} maxs {maxs}")                       # runner.rs:273-275
    sc = FdtdSolverConfig(fdtd.resolution, common.physical_constants, tuple(float(v) for v in size))
    lattice = sc.lattice_size()
    courant = min(fdtd.resolution.spatial) / (common.physical_constants.speed_of_light() * math.sqrt(3.0))   # fdtd/mod.rs:40-45
    courant_ok = fdtd.resolution.temporal <= courant
    if not courant_ok:
        warnings.warn(f"resolution doesn't satisfy courant condition (dt {fdtd.resolution.temporal} > {courant:.6g})")   # runner.rs:283-291

    if device is None:
        device = common.parallelization.device if common.parallelization is not None and common.parallelization.kind == "B200" else 0
    backend = FdtdB200Backend(device=device, exact_arithmetic=exact_arithmetic, memory_limit=common.memory_limit)
    need = backend.memory_required(sc) or 0
    if common.memory_limit is not None and need > common.memory_limit:                          # runner.rs:304-330
        raise MemoryError(f"lattice {lattice} needs {need} bytes of device memory, memory_limit is {common.memory_limit}")

    tr = CoordinateTransformations.for_fdtd(fdtd.resolution, lattice, common.volume.rotation(), mins)
    shapes = [_shape(e) for e in cfg.scene if e.collider is not None and e.material is not None and e.pml is None]
    regions = [PmlRegion(tuple(e.collider[1]), e.pml, e.transform.translation, e.transform.rotation)
               for e in cfg.scene if e.pml is not None and e.collider is not None and e.collider[0] == "Cuboid"]
    domain = SceneDomain(shapes, tr, common.default_material, pml_regions=regions, resolution=fdtd.resolution,
                         physical_constants=common.physical_constants)
    instance = backend.create_instance(sc, domain)
    state = instance.create_state()

    sources, skipped = Sources(), 0
    for e in cfg.scene:                                                                          # runner.rs:852-869
        if e.source is None:
            continue
        p = tr.transform_point_from_world_to_solver(e.transform.translation)
        if p is None:
            skipped += 1
            continue
        sources.push(p, e.source)

    observers = []
    k = 0
    for e in cfg.scene:                                                                          # runner.rs:732-818
        if e.observer is None:
            continue
        spec = e.observer
        size_xy = (lattice[0], lattice[1])
        if spec.write_to_gif:
            target = fio.GifEncoder(_join(out_dir, spec.write_to_gif), size_xy, frame_delay_ms=int(1000 * (cfg.observation_delay or 0.04)))
        elif spec.write_to_png:
            target = fio.PngSequence(_join(out_dir, spec.write_to_png), size_xy)
        elif out_dir:
            target = fio.PngSequence(_join(out_dir, f"observer{k}_{{:05d}}.png"), size_xy)
        else:
            target = ImageTarget(*size_xy)
        cmap = np.eye(4, dtype=np.float32) if spec.color_map is None else np.asarray(spec.color_map, dtype=np.float32).reshape(4, 4)
        params = ProjectionParameters(np.eye(4, dtype=np.float32), FieldComponent.E if spec.field == "E" else FieldComponent.H, cmap,
                                      APP_COLOR_MAP_CODE if spec.app_color_map else None)
        observers.append(Observer(spec, instance.create_projection(state, target, params)))
        k += 1
    run = PreparedRun(cfg, sc, lattice, tr, instance, state, sources, observers, int(need), skipped, courant_ok)
    run.timings["prepare_s"] = _time.monotonic() - t0
    return run


def _join(out_dir: Optional[str], name: str) -> str:
    import os

    if out_dir and not os.path.isabs(name):
        os.makedirs(out_dir, exist_ok=True)
        return os.path.join(out_dir, name)
    return name


def observe(run: PreparedRun) -> None:
    """`Observers::run` (runner.rs:710-728): one projection pass over every observer; errors stop the run (runner.rs:536-540)."""
    if not run.observers:
        return
    p = run.instance.begin_projection_pass(run.state)
    for o in run.observers:
        p.add_projection(o.projection)
        o.frames += 1
    p.finish()


def run(prepared: PreparedRun, stop: Optional[StopCondition] = None, batched: bool = False, observe_every: int = 0,
        on_step: Optional[Callable[[PreparedRun], None]] = None) -> dict:
    """The solver thread's loop (runner.rs:471-557).  Returns {steps, seconds, observations, ...}.

    Reference protocol (default): per step `begin_update → sources.apply(t) → finish`, observers when more than
    `observation_delay` seconds of wall clock have passed since the last observation (and once before the first step,
    runner.rs:481), `step_delay` honoured by sleeping.  `batched`: the steps up to the next observation (`observe_every`, or the
    stop condition's limit) are one `fdtd_b200_step(h, n)` with a device-resident forcing table."""
    cfg = prepared.config
    stop = stop or cfg.fdtd.stop_condition
    st, inst = prepared.state, prepared.instance
    dt = cfg.fdtd.resolution.temporal
    t_start = _time.monotonic()
    steps = observations = 0
    observe(prepared)                                                   # runner.rs:481
    observations += bool(prepared.observers)
    last_obs = _time.monotonic()
    if batched:
        if stop.kind == "never":
            raise ValueError("a batched run needs a step or simulated-time limit")
        while not stop.reached(_time.monotonic() - t_start, st):
            if stop.kind == "step_limit":
                left = int(stop.limit) - st.tick()
            elif stop.kind == "simulated_time_limit":
                left = max(int(math.ceil((float(stop.limit) - st.time()) / dt)), 1)
            else:
                left = observe_every or 256
            n = max(min(left, observe_every or left, 4096), 1)
            if prepared.sources.sources:
                st.set_source_table([p for p, _ in prepared.sources.sources], prepared.sources.table(st.time(), dt, n))
            st.step(n)
            steps += n
            if prepared.observers:
                observe(prepared)
                observations += 1
            if on_step is not None:
                on_step(prepared)
        st.sync()
    else:
        while not stop.reached(_time.monotonic() - t_start, st):
            t_pass = _time.monotonic()
            sim_time = st.time()
            up = inst.begin_update(st)
            prepared.sources.apply(sim_time, up)
            up.finish()
            steps += 1
            due = (observe_every and steps % observe_every == 0) if observe_every else (
                cfg.observation_delay is not None and _time.monotonic() - last_obs > cfg.observation_delay)
            if due and prepared.observers:
                observe(prepared)
                observations += 1
                last_obs = _time.monotonic()
            if on_step is not None:
                on_step(prepared)
            if cfg.step_delay:
                sleep = cfg.step_delay - (_time.monotonic() - t_pass)
                if sleep > 0:
                    _time.sleep(sleep)
        st.sync()
    secs = _time.monotonic() - t_start
    cells = prepared.lattice[0] * prepared.lattice[1] * prepared.lattice[2]
    return {"label": cfg.label, "lattice": list(prepared.lattice), "steps": steps, "tick": st.tick(), "time": st.time(), "seconds": secs,
            "observations": observations, "gcells_per_s": cells * steps / max(secs, 1e-12) / 1e9, "step_path": st.step_path,
            "memory_required": prepared.memory_required, "memory_limit": cfg.common.memory_limit, "sources": len(prepared.sources.sources),
            "sources_outside_lattice": prepared.skipped_sources, "courant_ok": prepared.courant_ok, "protocol": "batched" if batched else "per-step"}


def main(argv=None) -> int:
    ap = argparse.ArgumentParser(prog="python -m fdtd_b200.run", description=__doc__.split("\n\n")[0])
    ap.add_argument("config", help="SolverConfig file (.json or .toml); `app` = the application's canned test config (composer/mod.rs:841-869)")
    ap.add_argument("--steps", type=int, default=0, help="override the stop condition with StepLimit{steps}")
    ap.add_argument("--batched", action="store_true", help="device-resident forcing, many steps per call")
    ap.add_argument("--observe-every", type=int, default=0, help="observe every N steps instead of every observation_delay seconds")
    ap.add_argument("--out", default=None, help="directory for observer frames (PNG sequence unless the observer names a file)")
    ap.add_argument("--exact", action="store_true", help="FDTD_B200_EXACT_ARITHMETIC")
    ap.add_argument("--device", type=int, default=None)
    ap.add_argument("--dump-default-config", action="store_true", help="print the canned config as JSON and exit")
    args = ap.parse_args(argv)
    if args.dump_default_config:
        print(json.dumps(cfgmod.APP_TEST_CONFIG, indent=2))
        return 0
    cfg = cfgmod.from_dict(cfgmod.APP_TEST_CONFIG) if args.config == "app" else cfgmod.load(args.config)
    stop = StopCondition("step_limit", args.steps) if args.steps > 0 else cfg.fdtd.stop_condition
    if stop.kind == "never":
        print("stop_condition is Never and no --steps given: nothing would end this headless run", file=sys.stderr)
        return 2
    try:
        prepared = prepare(cfg, out_dir=args.out, exact_arithmetic=args.exact, device=args.device)
    except MemoryError as e:
        print(f"error: {e}", file=sys.stderr)     # the reference bails here too (runner.rs:319-330)
        return 1
    try:
        print(json.dumps(run(prepared, stop, batched=args.batched, observe_every=args.observe_every)))
    finally:
        prepared.close()
    return 0


if __name__ == "__main__":
    sys.exit(main())
