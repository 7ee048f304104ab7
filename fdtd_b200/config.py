"""`SolverConfig` files: the reference's solver configuration, loaded from JSON or TOML.

Shape and field names follow cem-app/src/solver/config.rs:27-171 as serde writes them (externally tagged enums):

    {"label": "...",
     "common":    {"volume": {"Fixed": {"isometry": {"rotation": [i, j, k, w], "translation": [x, y, z]}, "half_extents": [..]}}
                            | {"SceneAabb": {"rotation": [i, j, k, w], "margin": [..]}},
                   "physical_constants": {"vacuum_permittivity": 1.0, "vacuum_permeability": 1.0},
                   "default_material": {"relative_permeability": ..., "magnetic_conductivity": ..., "relative_permittivity": ...,
                                        "eletrical_conductivity": ...},
                   "parallelization": null | "Wgpu" | {"MultiThreaded": {"num_threads": null}} | {"B200": {"device": 0}},
                   "memory_limit": null | bytes},
     "specifics": {"Fdtd": {"resolution": {"spatial": [dx, dy, dz], "temporal": dt},
                            "stop_condition": "Never" | {"StepLimit": {"limit": n}} | {"SimulatedTimeLimit": {"limit": t}}
                                              | {"RealtimeLimit": {"limit": {"secs": s, "nanos": n}}}}}}

The reference keeps its scene in a bevy_ecs world (RON project files, cem-app/src/composer/file_formats); a headless run
needs one too, so a config file may carry a `"scene"` list next to the reference's three keys — entities with a transform and
any of a collider + material, a source, an observer, a graded PML (the components `cem-app/src/solver/runner.rs` queries:
:606-631 materials, :852-869 sources, :732-818 observers, :594-604 PMLs).  `Parallelization::B200` is the variant the Rust shim
adds (rust/cem-app-patch/runner_b200.rs); every other variant is accepted and runs on the B200 engine all the same — this
package has no CPU or wgpu path.
"""
from __future__ import annotations

import json
import math
from dataclasses import dataclass, field
from typing import Optional, Sequence

import numpy as np

from .driver import StopCondition
from .solver import (ContinousWave, FdtdSolverConfig, GaussianPulse, GradedPml, Material, PhysicalConstants, Resolution)


class ConfigError(ValueError):
    pass


def _need(d: dict, key: str, where: str):
    if not isinstance(d, dict) or key not in d:
        raise ConfigError(f"{where}: missing `{key}`")
    return d[key]


def _vec(v, n: int, where: str) -> tuple:
    if not isinstance(v, (list, tuple)) or len(v) != n:
        raise ConfigError(f"{where}: expected {n} numbers, got {v!r}")
    return tuple(float(x) for x in v)


def _variant(v, where: str) -> tuple:
    """serde's externally tagged enum: "Unit" or {"Variant": body}."""
    if isinstance(v, str):
        return v, None
    if isinstance(v, dict) and len(v) == 1:
        (k, body), = v.items()
        return k, body
    raise ConfigError(f"{where}: expected an enum variant, got {v!r}")


def _quat_wxyz(q_ijkw: Sequence[float]) -> tuple:
    """nalgebra serialises a UnitQuaternion as its coordinate vector [i, j, k, w]; scene.py wants (w, x, y, z)."""
    i, j, k, w = q_ijkw
    return (w, i, j, k)


@dataclass
class Isometry:
    translation: tuple = (0.0, 0.0, 0.0)
    rotation: tuple = (1.0, 0.0, 0.0, 0.0)     # w, x, y, z

    @staticmethod
    def parse(d, where: str) -> "Isometry":
        if d is None:
            return Isometry()
        rot = d.get("rotation", [0.0, 0.0, 0.0, 1.0])
        return Isometry(_vec(d.get("translation", [0.0, 0.0, 0.0]), 3, where + ".translation"), _quat_wxyz(_vec(rot, 4, where + ".rotation")))


@dataclass
class Volume:
    """config.rs:98-171.  kind: "Fixed" (isometry + half_extents) | "SceneAabb" (rotation + margin)."""
    kind: str = "SceneAabb"
    isometry: Isometry = field(default_factory=Isometry)
    half_extents: tuple = (1.0, 1.0, 1.0)
    margin: tuple = (0.0, 0.0, 0.0)

    @staticmethod
    def parse(v, where: str) -> "Volume":
        kind, body = _variant(v, where)
        body = body or {}
        if kind == "Fixed":
            return Volume("Fixed", Isometry.parse(body.get("isometry"), where + ".isometry"), _vec(_need(body, "half_extents", where), 3, where + ".half_extents"))
        if kind == "SceneAabb":
            rot = _quat_wxyz(_vec(body.get("rotation", [0.0, 0.0, 0.0, 1.0]), 4, where + ".rotation"))
            return Volume("SceneAabb", Isometry((0.0, 0.0, 0.0), rot), margin=_vec(body.get("margin", [0.0, 0.0, 0.0]), 3, where + ".margin"))
        raise ConfigError(f"{where}: unknown Volume variant {kind!r}")

    def rotation(self) -> tuple:
        return self.isometry.rotation


@dataclass
class Parallelization:
    """config.rs:83-87 plus the shim's variant."""
    kind: str = "B200"                  # MultiThreaded | Wgpu | B200
    num_threads: Optional[int] = None
    device: int = 0

    @staticmethod
    def parse(v, where: str) -> Optional["Parallelization"]:
        if v is None:
            return None
        kind, body = _variant(v, where)
        body = body or {}
        if kind == "MultiThreaded":
            return Parallelization("MultiThreaded", num_threads=body.get("num_threads"))
        if kind == "Wgpu":
            return Parallelization("Wgpu")
        if kind == "B200":
            return Parallelization("B200", device=int(body.get("device", 0)))
        raise ConfigError(f"{where}: unknown Parallelization variant {kind!r}")


@dataclass
class SolverConfigCommon:
    volume: Volume
    physical_constants: PhysicalConstants
    default_material: Material
    parallelization: Optional[Parallelization] = None
    memory_limit: Optional[int] = None


@dataclass
class SolverConfigFdtd:
    resolution: Resolution
    stop_condition: StopCondition


@dataclass
class SolverConfig:
    label: str
    common: SolverConfigCommon
    fdtd: SolverConfigFdtd
    scene: list = field(default_factory=list)   # [SceneEntity]
    observation_delay: Optional[float] = 0.04    # seconds; runner.rs:464 (25 observations per second)
    step_delay: Optional[float] = None           # seconds; the reference's UI default is 10 ms (runner.rs:463), a headless run has none


def _material(d, where: str) -> Material:
    if d is None:
        return Material.VACUUM
    known = ("relative_permeability", "magnetic_conductivity", "relative_permittivity", "eletrical_conductivity")
    for k in d:
        if k not in known:
            raise ConfigError(f"{where}: unknown Material field {k!r} (the reference spells it `eletrical_conductivity`)")
    v = Material.VACUUM
    return Material(float(d.get(known[0], v.relative_permeability)), float(d.get(known[1], v.magnetic_conductivity)),
                    float(d.get(known[2], v.relative_permittivity)), float(d.get(known[3], v.eletrical_conductivity)))


def _constants(v, where: str) -> PhysicalConstants:
    if v is None or v == "SI":
        return PhysicalConstants.SI
    if v == "REDUCED":
        return PhysicalConstants.REDUCED
    return PhysicalConstants(float(_need(v, "vacuum_permittivity", where)), float(_need(v, "vacuum_permeability", where)))


def _stop_condition(v, where: str) -> StopCondition:
    kind, body = _variant(v, where)
    if kind == "Never":
        return StopCondition("never")
    limit = _need(body or {}, "limit", where)
    if kind == "StepLimit":
        return StopCondition("step_limit", int(limit))
    if kind == "SimulatedTimeLimit":
        return StopCondition("simulated_time_limit", float(limit))
    if kind == "RealtimeLimit":      # std::time::Duration → {"secs": .., "nanos": ..}
        secs = float(limit.get("secs", 0)) + 1e-9 * float(limit.get("nanos", 0)) if isinstance(limit, dict) else float(limit)
        return StopCondition("realtime_limit", secs)
    raise ConfigError(f"{where}: unknown StopCondition variant {kind!r}")


# ---- scene entities (what the runner's ECS queries would find) ----------------------------------------------------------------------

@dataclass
class ObserverSpec:
    """cem-app/src/solver/observer.rs:30-60: which field, the linear colour map, and where the frames go."""
    field: str = "E"
    color_map: Optional[list] = None        # 4x4 row-major; None = identity
    app_color_map: bool = True              # the application's value.z red/green snippet (runner.rs:761-780)
    write_to_gif: Optional[str] = None
    write_to_png: Optional[str] = None      # pattern with {} for the frame number


@dataclass
class SceneEntity:
    transform: Isometry = field(default_factory=Isometry)
    collider: Optional[tuple] = None        # ("Ball", radius) | ("Cuboid", (hx, hy, hz)) | ("Cylinder", half_height, radius)
    material: Optional[Material] = None
    source: Optional[object] = None         # solver.WithAmplitudes
    observer: Optional[ObserverSpec] = None
    pml: Optional[GradedPml] = None


def _collider(v, where: str) -> tuple:
    kind, body = _variant(v, where)
    body = body or {}
    if kind == "Ball":
        return ("Ball", float(_need(body, "radius", where)))
    if kind == "Cuboid":
        return ("Cuboid", _vec(_need(body, "half_extents", where), 3, where + ".half_extents"))
    if kind == "Cylinder":
        return ("Cylinder", float(_need(body, "half_height", where)), float(_need(body, "radius", where)))
    raise ConfigError(f"{where}: unsupported collider {kind!r} (Ball, Cuboid, Cylinder)")


def _source(d, where: str):
    """{"GaussianPulse": {"time": t0, "duration": tau}} | {"ContinousWave": {"phase": p, "frequency": f}} with "j" and "m"
    amplitudes next to it (cem-solver/src/source.rs:39-98)."""
    fn = None
    for kind in ("GaussianPulse", "ContinousWave"):
        if kind in d:
            b = d[kind]
            fn = (GaussianPulse(float(_need(b, "time", where)), float(_need(b, "duration", where))) if kind == "GaussianPulse"
                  else ContinousWave(float(b.get("phase", 0.0)), float(_need(b, "frequency", where))))
    if fn is None:
        raise ConfigError(f"{where}: a source needs a GaussianPulse or a ContinousWave")
    return fn.with_amplitudes(_vec(d.get("j", [0, 0, 0]), 3, where + ".j"), _vec(d.get("m", [0, 0, 0]), 3, where + ".m"))


def _entity(d, where: str) -> SceneEntity:
    e = SceneEntity(transform=Isometry.parse(d.get("transform"), where + ".transform"))
    if "collider" in d:
        e.collider = _collider(d["collider"], where + ".collider")
    if "material" in d:
        e.material = _material(d["material"], where + ".material")
    if "source" in d:
        e.source = _source(d["source"], where + ".source")
    if "observer" in d:
        o = d["observer"] or {}
        e.observer = ObserverSpec(str(o.get("field", "E")), o.get("color_map"), bool(o.get("app_color_map", True)), o.get("write_to_gif"), o.get("write_to_png"))
        if e.observer.field not in ("E", "H"):
            raise ConfigError(f"{where}.observer.field must be E or H")
    if "pml" in d:
        p = d["pml"]
        g = GradedPml()
        e.pml = GradedPml(float(p.get("m", g.m)), float(p.get("m_a", g.m_a)), float(p.get("sigma_max", g.sigma_max)), float(p.get("kappa_max", g.kappa_max)),
                          float(p.get("a_max", g.a_max)), _vec(p.get("normal", [1, 0, 0]), 3, where + ".pml.normal"))
    return e


def from_dict(d: dict) -> SolverConfig:
    common = _need(d, "common", "config")
    kind, body = _variant(_need(d, "specifics", "config"), "specifics")
    if kind != "Fdtd":
        raise ConfigError(f"specifics: only Fdtd runs (the reference's Feec arm logs `todo: feec solver`, runner.rs:157); got {kind!r}")
    res = _need(body, "resolution", "specifics.Fdtd")
    limit = common.get("memory_limit")
    cfg = SolverConfig(
        label=str(d.get("label", "")),
        common=SolverConfigCommon(Volume.parse(common.get("volume", "SceneAabb"), "common.volume"),
                                  _constants(common.get("physical_constants"), "common.physical_constants"),
                                  _material(common.get("default_material"), "common.default_material"),
                                  Parallelization.parse(common.get("parallelization"), "common.parallelization"),
                                  None if limit is None else int(limit)),
        fdtd=SolverConfigFdtd(Resolution(_vec(_need(res, "spatial", "resolution"), 3, "resolution.spatial"), float(_need(res, "temporal", "resolution"))),
                              _stop_condition(body.get("stop_condition", "Never"), "specifics.Fdtd.stop_condition")),
        scene=[_entity(e, f"scene[{k}]") for k, e in enumerate(d.get("scene", []))])
    if "observation_delay" in d:
        cfg.observation_delay = None if d["observation_delay"] is None else float(d["observation_delay"])
    if "step_delay" in d:
        cfg.step_delay = None if d["step_delay"] is None else float(d["step_delay"])
    return cfg


def loads(text: str, fmt: str = "json") -> SolverConfig:
    if fmt == "toml":
        import tomllib

        return from_dict(tomllib.loads(text))
    return from_dict(json.loads(text))


def load(path: str) -> SolverConfig:
    """JSON (`.json`) or TOML (`.toml`), the two formats the application's own config files use (cem-app/src/config.rs:17-35)."""
    with open(path, "r", encoding="utf-8") as f:
        return loads(f.read(), "toml" if path.lower().endswith(".toml") else "json")


APP_TEST_CONFIG = {
    # cem-app/src/composer/mod.rs:841-869 `make_config`, restated as the JSON serde would write (plus the example scene's source,
    # presets.rs:176-178, and one observer): volume 1 x 1 x 0 at dx = 0.01 -> a 100 x 100 x 1 lattice
    "label": "Test FDTD (B200)",
    "common": {
        "volume": {"Fixed": {"isometry": {"rotation": [0.0, 0.0, 0.0, 1.0], "translation": [0.0, 0.5, 0.0]}, "half_extents": [0.5, 0.5, 0.0]}},
        "physical_constants": {"vacuum_permittivity": 1.0, "vacuum_permeability": 1.0},
        "default_material": {"relative_permeability": 1.0, "magnetic_conductivity": 0.0, "relative_permittivity": 1.0, "eletrical_conductivity": 10.0},
        "parallelization": {"B200": {"device": 0}},
        "memory_limit": 200000000,
    },
    "specifics": {"Fdtd": {"resolution": {"spatial": [0.01, 0.01, 0.01], "temporal": 0.001}, "stop_condition": "Never"}},
    "scene": [
        {"transform": {"translation": [0.0, 0.5, 0.0]}, "source": {"GaussianPulse": {"time": 0.05, "duration": 0.01}, "j": [0.0, 0.0, 50.0], "m": [0.0, 0.0, 0.0]}},
        {"transform": {"translation": [0.2, 0.7, 0.0]}, "collider": {"Ball": {"radius": 0.1}}, "material": {"relative_permittivity": 3.9}},
        {"observer": {"field": "E"}},
    ],
}
