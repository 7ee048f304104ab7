"""Headless solver driver: the loop of cem-app/src/solver/runner.rs:471-557 without the UI thread, step delay
and pause machinery.  Two ways to run:
  * `run_reference_style` — begin_update / sources.apply(t) / finish per step, exactly as the reference's
    solver thread calls a backend (one host round-trip per step);
  * `run_batched` — sources pre-tabulated on the host in f64 with the same `time += dt` accumulation, then
    one fdtd_b200_step(n) (device-resident forcing table, no host round-trips).
Both produce identical fields (tests/test_parity_gpu.py)."""
from __future__ import annotations

import time as _time
from dataclasses import dataclass
from typing import Callable, Optional, Sequence

import numpy as np

from .solver import FdtdB200SolverInstance, FdtdB200SolverState


@dataclass
class StopCondition:
    """cem-app/src/solver/config.rs StopCondition + runner.rs:933-947."""
    kind: str = "never"        # never | step_limit | simulated_time_limit | realtime_limit
    limit: float = 0.0

    def reached(self, elapsed_s: float, state: FdtdB200SolverState) -> bool:
        if self.kind == "never":
            return False
        if self.kind == "step_limit":
            return state.tick() >= int(self.limit)
        if self.kind == "simulated_time_limit":
            return np.float32(state.time()) >= np.float32(self.limit)  # `state.time() as f32 >= limit`
        if self.kind == "realtime_limit":
            return elapsed_s >= self.limit
        raise ValueError(self.kind)


class Sources:
    """cem-app/src/solver/runner.rs:821-849."""

    def __init__(self, sources: Sequence[tuple] = ()):
        self.sources = list(sources)  # [(point, source with .evaluate(t) -> SourceValues)]

    def push(self, point, source) -> None:
        self.sources.append((point, source))

    def apply(self, time: float, update_pass) -> None:
        for point, source in self.sources:
            update_pass.set_forcing(point, source.evaluate(time))

    def table(self, t0: float, dt: float, n_steps: int) -> np.ndarray:
        out = np.zeros((n_steps, len(self.sources), 6), dtype=np.float64)
        t = float(t0)
        for s in range(n_steps):
            for k, (_, src) in enumerate(self.sources):
                v = src.evaluate(t)
                out[s, k, :3] = v.j
                out[s, k, 3:] = v.m
            t += dt  # accumulated like cpu/mod.rs:513
        return out


def run_reference_style(instance: FdtdB200SolverInstance, state: FdtdB200SolverState, sources: Sources, stop: StopCondition,
                        observe: Optional[Callable[[FdtdB200SolverState], None]] = None, observe_every: int = 0) -> int:
    """runner.rs:512-542.  Returns the number of steps taken."""
    t_start = _time.monotonic()
    steps = 0
    while not stop.reached(_time.monotonic() - t_start, state):
        sim_time = state.time()
        update_pass = instance.begin_update(state)
        sources.apply(sim_time, update_pass)
        update_pass.finish()
        steps += 1
        if observe is not None and observe_every and steps % observe_every == 0:
            observe(state)
    return steps


def run_batched(state: FdtdB200SolverState, sources: Sources, n_steps: int, dt: float, chunk: int = 4096) -> None:
    """Same physics, device-resident forcing: tabulate in f64 on the host, cast once, step(n)."""
    done = 0
    t = state.time()
    while done < n_steps:
        n = min(chunk, n_steps - done)
        if sources.sources:
            tab = sources.table(t, dt, n)
            state.set_source_table([p for p, _ in sources.sources], tab)
        state.step(n)
        for _ in range(n):
            t += dt
        done += n
