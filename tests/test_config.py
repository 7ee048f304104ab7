"""Config-file driven runs: the reference's `SolverConfig` (cem-app/src/solver/config.rs:27-171) from JSON and
TOML, the application's canned test configuration (cem-app/src/composer/mod.rs:841-869) restated as such a file, and the headless
runner built on it (fdtd_b200/run.py ≙ cem-app/src/solver/runner.rs:247-380, 471-557)."""
import json

import numpy as np
import pytest

from helpers import oracle_run

APP_TOML = """
label = "Test FDTD (toml)"

[common]
memory_limit = 200000000
parallelization = "Wgpu"

[common.volume.Fixed]
half_extents = [0.5, 0.5, 0.0]
[common.volume.Fixed.isometry]
rotation = [0.0, 0.0, 0.0, 1.0]
translation = [0.0, 0.5, 0.0]

[common.physical_constants]
vacuum_permittivity = 1.0
vacuum_permeability = 1.0

[common.default_material]
relative_permeability = 1.0
magnetic_conductivity = 0.0
relative_permittivity = 1.0
eletrical_conductivity = 10.0

[specifics.Fdtd]
stop_condition = { StepLimit = { limit = 25 } }
[specifics.Fdtd.resolution]
spatial = [0.01, 0.01, 0.01]
temporal = 0.001
"""


def test_the_apps_canned_config_loads_from_json_and_toml(b200, tmp_path):
    from fdtd_b200 import config, run

    path = tmp_path / "app.json"
    path.write_text(json.dumps(config.APP_TEST_CONFIG))
    cfg = config.load(str(path))
    tpath = tmp_path / "app.toml"
    tpath.write_text(APP_TOML)
    tcfg = config.load(str(tpath))
    for c in (cfg, tcfg):
        assert c.common.volume.kind == "Fixed" and c.common.volume.half_extents == (0.5, 0.5, 0.0)
        assert c.common.volume.isometry.translation == (0.0, 0.5, 0.0) and c.common.volume.rotation() == (1.0, 0.0, 0.0, 0.0)
        assert c.common.physical_constants == b200.PhysicalConstants.REDUCED
        assert c.common.default_material == b200.Material(eletrical_conductivity=10.0)       # "intoduce dissipation", mod.rs:853-856
        assert c.common.memory_limit == 200_000_000
        assert c.fdtd.resolution.spatial == (0.01, 0.01, 0.01) and c.fdtd.resolution.temporal == 0.001
        mins, maxs = run.volume_aabb(c.common.volume, c.scene)
        sc = b200.FdtdSolverConfig(c.fdtd.resolution, c.common.physical_constants, tuple(float(v) for v in (maxs - mins)))
        assert sc.lattice_size() == (100, 100, 1)                                            # the only lattice the application ships
        be = b200.FdtdB200Backend(memory_limit=c.common.memory_limit)
        assert be.memory_required(sc) == 12 * 4 * 128 * 100 * 3 + 128 * 100 * 2 + 16 < c.common.memory_limit
    assert cfg.common.parallelization.kind == "B200" and cfg.fdtd.stop_condition.kind == "never" and len(cfg.scene) == 3
    assert tcfg.common.parallelization.kind == "Wgpu" and tcfg.fdtd.stop_condition.kind == "step_limit" and tcfg.fdtd.stop_condition.limit == 25
    assert cfg.scene[0].source is not None and cfg.scene[1].collider == ("Ball", 0.1) and cfg.scene[2].observer.field == "E"
    assert cfg.observation_delay == 0.04 and cfg.step_delay is None                          # runner.rs:464; a headless run is not UPS limited


def test_enum_variants_and_errors():
    from fdtd_b200 import config

    base = json.loads(json.dumps(config.APP_TEST_CONFIG))
    for sc, kind, limit in (({"StepLimit": {"limit": 7}}, "step_limit", 7), ({"SimulatedTimeLimit": {"limit": 0.5}}, "simulated_time_limit", 0.5),
                            ({"RealtimeLimit": {"limit": {"secs": 2, "nanos": 500000000}}}, "realtime_limit", 2.5), ("Never", "never", 0.0)):
        base["specifics"]["Fdtd"]["stop_condition"] = sc
        c = config.from_dict(base)
        assert c.fdtd.stop_condition.kind == kind and c.fdtd.stop_condition.limit == limit
    for par, kind in ((None, None), ("Wgpu", "Wgpu"), ({"MultiThreaded": {"num_threads": None}}, "MultiThreaded"), ({"B200": {"device": 3}}, "B200")):
        base["common"]["parallelization"] = par
        p = config.from_dict(base).common.parallelization
        assert (p is None) if kind is None else p.kind == kind
    base["common"]["volume"] = {"SceneAabb": {"rotation": [0, 0, 0, 1], "margin": [0.1, 0.1, 0.0]}}
    assert config.from_dict(base).common.volume.kind == "SceneAabb"
    bad = json.loads(json.dumps(config.APP_TEST_CONFIG))
    bad["specifics"] = {"Feec": {}}
    with pytest.raises(config.ConfigError, match="only Fdtd"):
        config.from_dict(bad)
    bad = json.loads(json.dumps(config.APP_TEST_CONFIG))
    bad["common"]["default_material"]["electrical_conductivity"] = 1.0                       # the reference's spelling is `eletrical_`
    with pytest.raises(config.ConfigError, match="eletrical_conductivity"):
        config.from_dict(bad)
    bad = json.loads(json.dumps(config.APP_TEST_CONFIG))
    del bad["specifics"]["Fdtd"]["resolution"]["temporal"]
    with pytest.raises(config.ConfigError, match="temporal"):
        config.from_dict(bad)


def test_scene_aabb_volume_and_memory_limit_bail(b200):
    from fdtd_b200 import config, run

    d = json.loads(json.dumps(config.APP_TEST_CONFIG))
    d["common"]["volume"] = {"SceneAabb": {"rotation": [0, 0, 0, 1], "margin": [0.05, 0.05, 0.0]}}
    cfg = config.from_dict(d)
    mins, maxs = run.volume_aabb(cfg.common.volume, cfg.scene)       # the ball of radius 0.1 at (0.2, 0.7, 0) plus the margin
    assert np.allclose(mins, [0.05, 0.55, -0.1]) and np.allclose(maxs, [0.35, 0.85, 0.1])
    d = json.loads(json.dumps(config.APP_TEST_CONFIG))
    d["common"]["memory_limit"] = 1000
    with pytest.raises(MemoryError, match="memory_limit is 1000"):   # before anything touches a device (runner.rs:304-330)
        run.prepare(config.from_dict(d))


@pytest.mark.gpu
def test_the_apps_config_runs_against_the_oracle(b200, oracle, tmp_path, monkeypatch):
    """`python -m fdtd_b200.run app --steps 60`: per-step protocol with the wall-clock observation cadence, and the batched
    protocol — both bit-identical to the fp32 oracle on the lattice the scene rasterises to; frames land in the PNG sequence."""
    from fdtd_b200 import config, configs, run
    from fdtd_b200 import io as fio

    monkeypatch.setenv("FDTD_B200_RESIDENT", "1")
    steps = 60
    cfg = config.from_dict(config.APP_TEST_CONFIG)
    cfg.observation_delay = 0.0                                       # every step is "due": exercises the cadence branch
    results = {}
    for batched in (False, True):
        p = run.prepare(cfg, out_dir=str(tmp_path / ("b" if batched else "s")), exact_arithmetic=True)
        assert p.lattice == (100, 100, 1) and p.sources.sources[0][0] == (50, 50, 0) and p.memory_required < 200_000_000
        out = run.run(p, b200_stop(steps), batched=batched, observe_every=20 if batched else 0)
        assert out["steps"] == steps and out["tick"] == steps and "resident" in out["step_path"]
        assert out["observations"] == (1 + 3 if batched else 1 + steps)
        ids = p.state.material_ids()
        e, h = p.state.read_field(b200.FieldComponent.E), p.state.read_field(b200.FieldComponent.H)
        results[batched] = (e, h)
        pal = p.instance.palette
        p.close()
    assert np.array_equal(results[False][0], results[True][0]) and np.array_equal(results[False][1], results[True][1])
    assert len(pal) == 2 and 0 < int((ids == 1).sum()) < 400          # the eps_r = 3.9 ball: ~314 cells
    w = configs.Workload("app", (100, 100, 1), cfg.fdtd.resolution, cfg.common.physical_constants, pal, ids, [((50, 50, 0), cfg.scene[0].source)], [], 0)
    ref, _ = oracle_run(oracle, w, steps, oracle.F32)
    assert np.array_equal(results[True][0], ref.field(oracle.E).astype(np.float32)) and np.any(results[True][0])
    assert np.array_equal(results[True][1], ref.field(oracle.H).astype(np.float32))
    import glob
    assert len(glob.glob(str(tmp_path / "s" / "observer0_*.png"))) == 1 + steps


def b200_stop(n):
    from fdtd_b200.driver import StopCondition

    return StopCondition("step_limit", n)
