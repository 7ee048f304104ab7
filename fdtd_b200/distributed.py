"""z-slab decomposition of the lattice over ranks (one process per GPU) and the one-plane E/H halo
exchange per half-step.  torch.distributed is plumbing only: NCCL send/recv between
z-neighbours, no collective on the data path.

Dependencies of the Yee update across a slab boundary (update.wgsl:147-177):
  H half-step of rank r needs Ex,Ey of plane z_begin-1      → after each E half-step every rank sends
                                                              its TOP Ex,Ey plane UP;
  E half-step of rank r needs the new Hx,Hy of plane z_end  → after each H half-step every rank sends
                                                              its BOTTOM Hx,Hy plane DOWN.
The boundary plane is updated first, its send is started, and the interior planes are updated while
the message is in flight.

The stepper is written against a small "slab engine" protocol (update_h, update_e, end_step, halo
tensors) so the orchestration is testable on CPU with gloo (tests/test_distributed_cpu.py).
"""
from __future__ import annotations

from typing import Optional, Protocol, Sequence

import numpy as np


def partition_z(nz: int, world: int) -> list:
    """Contiguous balanced z-slabs [(z_begin, z_end)] — earlier ranks take the remainder planes."""
    if world < 1 or nz < world:
        raise ValueError(f"cannot split {nz} planes over {world} ranks")
    base, rem = divmod(nz, world)
    out, z = [], 0
    for r in range(world):
        n = base + (1 if r < rem else 0)
        out.append((z, z + n))
        z += n
    return out


class SlabEngine(Protocol):
    n_planes: int

    def update_h(self, z_lo: int, z_hi: int) -> None: ...
    def update_e(self, z_lo: int, z_hi: int) -> None: ...
    def end_step(self) -> None: ...
    def halo(self, name: str):  # torch tensor; names: send_e_top, recv_e_low, send_h_bottom, recv_h_high
        ...


class CudaHaloView:
    """Exposes a device plane of the C library as a torch tensor without copying (__cuda_array_interface__)."""

    def __init__(self, ptr: int, n_floats: int):
        self.__cuda_array_interface__ = {"shape": (n_floats,), "typestr": "<f4", "data": (ptr, False), "version": 3, "strides": None}


class CudaSlabEngine:
    """SlabEngine over a FdtdB200SolverState that owns planes [z_begin, z_end) on this rank's GPU."""

    def __init__(self, state, device=None):
        import torch

        from . import _lib as L

        self.state = state
        self.n_planes = state.z_end - state.z_begin
        self._torch = torch
        dev = torch.device("cuda", torch.cuda.current_device()) if device is None else device
        state.set_stream(torch.cuda.current_stream(dev).cuda_stream)  # NCCL orders against torch's current stream

        def planes(a: int, b: int):
            out = []
            for which in (a, b):
                ptr, n = state.halo_plane(which)
                out.append(torch.as_tensor(CudaHaloView(ptr, n), device=dev))
            return out

        # two components per message (Ex,Ey / Hx,Hy) live in separate SoA arrays → two tensors each
        self._halo = {
            "send_e_top": planes(L.HALO_SEND_EX_TOP, L.HALO_SEND_EY_TOP),
            "recv_e_low": planes(L.HALO_RECV_EX_LOW, L.HALO_RECV_EY_LOW),
            "send_h_bottom": planes(L.HALO_SEND_HX_BOTTOM, L.HALO_SEND_HY_BOTTOM),
            "recv_h_high": planes(L.HALO_RECV_HX_HIGH, L.HALO_RECV_HY_HIGH),
        }
        # fused-step handles: the buffer sets alternate, so planes are looked up per call (tensors cached by address)
        self.fused = bool(state.uses_fused)
        self._dev = dev
        self._by_ptr = {}
        self._fused_names = {
            "send_up": (L.HALO_SEND_EX_TOP, L.HALO_SEND_EY_TOP),
            "recv_low": (L.HALO_RECV_EX_LOW, L.HALO_RECV_EY_LOW),
            "send_down": (L.HALO_SEND_EX_BOTTOM, L.HALO_SEND_EY_BOTTOM, L.HALO_SEND_EZ_BOTTOM, L.HALO_SEND_HX_BOTTOM, L.HALO_SEND_HY_BOTTOM),
            "recv_high": (L.HALO_RECV_EX_HIGH, L.HALO_RECV_EY_HIGH, L.HALO_RECV_EZ_HIGH, L.HALO_RECV_HX_HIGH, L.HALO_RECV_HY_HIGH),
            "send_mat_bottom": (L.HALO_SEND_MAT_BOTTOM,),
            "recv_mat_high": (L.HALO_RECV_MAT_HIGH,),
        }
        self._pending_bit = L.HALO_PENDING

    def update_h(self, z_lo: int, z_hi: int) -> None:
        self.state.update_h(z_lo, z_hi)

    def update_e(self, z_lo: int, z_hi: int) -> None:
        self.state.update_e(z_lo, z_hi)

    def end_step(self) -> None:
        self.state.end_step()

    def halo(self, name: str, pending: bool = False):
        if name in self._halo:
            return self._halo[name]
        out = []
        for which in self._fused_names[name]:
            ptr, n = self.state.halo_plane(which | (self._pending_bit if pending else 0))
            t = self._by_ptr.get((ptr, n))
            if t is None:
                t = self._torch.as_tensor(CudaHaloView(ptr, n), device=self._dev)
                self._by_ptr[(ptr, n)] = t
            out.append(t)
        return out

    def buffer_set_key(self) -> int:
        """Identifies the current ping-pong buffer set (address of one of its planes)."""
        return self.state.halo_plane(self._fused_names["send_up"][0])[0]

    def fused_update(self, z_lo: int, z_hi: int) -> None:
        self.state.fused_update(z_lo, z_hi)

    def fused_commit(self) -> None:
        self.state.fused_commit()


class SlabStepper:
    """Leapfrog step of one rank's slab with neighbour halo exchange over torch.distributed."""

    def __init__(self, engine: SlabEngine, rank: int, world: int, group=None, overlap: bool = True):
        self.engine, self.rank, self.world, self.group, self.overlap = engine, rank, world, group, overlap
        self.lower: Optional[int] = rank - 1 if rank > 0 else None
        self.upper: Optional[int] = rank + 1 if rank + 1 < world else None
        self.exchanges = 0
        self.fused = bool(getattr(engine, "fused", False))
        if self.fused:
            self._setup_fused()

    # -- transport -----------------------------------------------------------------------------
    def _start(self, send_name: str, send_to: Optional[int], recv_name: str, recv_from: Optional[int], **kw):
        import torch.distributed as dist

        ops = []
        # post receives first, lowest peer first on every rank → the pairing is deadlock-free for any backend
        if recv_from is not None:
            for t in self.engine.halo(recv_name, **kw):
                ops.append(dist.P2POp(dist.irecv, t, recv_from, self.group))
        if send_to is not None:
            for t in self.engine.halo(send_name, **kw):
                ops.append(dist.P2POp(dist.isend, t, send_to, self.group))
        if not ops:
            return []
        self.exchanges += 1
        return dist.batch_isend_irecv(ops)

    @staticmethod
    def _finish(works) -> None:
        for w in works:
            w.wait()

    def start_h_exchange(self):
        """After the H half-step: bottom Hx,Hy go DOWN, ghost-above comes from the upper rank."""
        return self._start("send_h_bottom", self.lower, "recv_h_high", self.upper)

    def start_e_exchange(self):
        """After the E half-step: top Ex,Ey go UP, ghost-below comes from the lower rank."""
        return self._start("send_e_top", self.upper, "recv_e_low", self.lower)

    def refresh_halos(self) -> None:
        """Make ghost planes consistent after fields were written from outside (write_field)."""
        if self.fused:
            self._finish(self._start_fused_exchange(pending=False))
            return
        self._finish(self.start_h_exchange())
        self._finish(self.start_e_exchange())

    # -- fused single-pass step: one exchange per step ---------------------------------------------
    def _setup_fused(self) -> None:
        """The fused kernel recomputes H' of the upper neighbour's first plane, so it needs that plane's material
        index (static, shipped once) besides its old E and Hx,Hy (shipped every step)."""
        self._fused_cache = {}
        self._finish(self._start("send_mat_bottom", self.lower, "recv_mat_high", self.upper))

    def _start_fused_exchange(self, pending: bool):
        """top Ex,Ey go UP (the upper rank's d/dz of its first plane); the whole first plane's E and Hx,Hy go DOWN
        (the lower rank recomputes H' there for its top plane's E').  One batched send/recv group per step; the
        tensor lists are cached per buffer-set parity (the two sets alternate) so the per-step host cost stays far
        below the 1.1 ms kernel."""
        import torch.distributed as dist

        key = (bool(pending), self.engine.buffer_set_key())   # which of the two ping-pong sets is current
        tensors = self._fused_cache.get(key)
        if tensors is None:
            tensors = []
            # receives first, lowest peer first on every rank → deadlock-free pairing for any backend
            for name, peer, recv in (("recv_low", self.lower, True), ("recv_high", self.upper, True),
                                     ("send_down", self.lower, False), ("send_up", self.upper, False)):
                if peer is not None:
                    tensors += [(recv, t, peer) for t in self.engine.halo(name, pending=pending)]
            self._fused_cache[key] = tensors
        if not tensors:
            return []
        self.exchanges += 1
        return dist.batch_isend_irecv([dist.P2POp(dist.irecv if recv else dist.isend, t, peer, self.group) for recv, t, peer in tensors])

    def _step_fused(self) -> None:
        e, n = self.engine, self.engine.n_planes
        lo = 1 if (self.lower is not None and self.overlap and n >= 3) else 0
        hi = n - 1 if (self.upper is not None and self.overlap and n >= 3) else n
        # boundary planes first (they are what the neighbours wait for), written into the pending buffer set
        if hi < n:
            e.fused_update(hi, n)
        if lo > 0:
            e.fused_update(0, lo)
        if lo > 0 or hi < n:
            works = self._start_fused_exchange(pending=True)   # overlaps the interior update below
            e.fused_update(lo, hi)
        else:
            e.fused_update(0, n)
            works = self._start_fused_exchange(pending=True)
        self._finish(works)
        e.fused_commit()
        e.end_step()

    # -- stepping ------------------------------------------------------------------------------
    def step(self, n_steps: int = 1) -> None:
        e, n = self.engine, self.engine.n_planes
        for _ in range(n_steps):
            if self.fused:
                self._step_fused()
                continue
            if self.overlap and n > 1:
                # H: the bottom plane feeds the lower rank — do it first, ship it, then the rest
                e.update_h(0, 1)
                works = self.start_h_exchange()
                e.update_h(1, n)
                self._finish(works)
                # E: the top plane needs the ghost H just received and feeds the upper rank
                e.update_e(n - 1, n)
                works = self.start_e_exchange()
                e.update_e(0, n - 1)
                self._finish(works)
            else:
                e.update_h(0, n)
                self._finish(self.start_h_exchange())
                e.update_e(0, n)
                self._finish(self.start_e_exchange())
            e.end_step()


class P2PSlabStepper:
    """Slab stepping over NVLink peer memory (fdtd_b200_p2p_*): torch.distributed only carries the 2 KB CUDA-IPC blobs
    between z-neighbours once; afterwards every step runs inside the C library — boundary planes on a high-priority stream,
    stored straight into the neighbours' ghost planes and ordered by device-side flags, interior planes concurrently on the
    main stream.  No NCCL call, no host synchronisation and no Python on the per-step path."""

    def __init__(self, state, rank: int, world: int, group=None):
        import torch
        import torch.distributed as dist

        self.state, self.rank, self.world, self.group = state, rank, world, group
        if not state.uses_fused:
            raise RuntimeError("the peer-memory transport needs a fused-step handle")
        blob = state.p2p_export()
        if world > 1:
            backend = dist.get_backend(group)
            dev = torch.device("cuda", torch.cuda.current_device()) if backend == "nccl" else torch.device("cpu")
            mine = torch.frombuffer(bytearray(blob), dtype=torch.uint8).to(dev)
            every = [torch.empty_like(mine) for _ in range(world)]
            dist.all_gather(every, mine, group=group)
            blobs = [bytes(t.cpu().numpy().tobytes()) for t in every]
        else:
            blobs = [blob]
        state.p2p_connect(blobs[rank - 1] if rank > 0 else None, blobs[rank + 1] if rank + 1 < world else None)
        if world > 1:
            dist.barrier(group=group)      # every rank has mapped its neighbours before anyone pushes
        self.refresh_halos()

    def refresh_halos(self) -> None:
        """After connect, reset or write_field: publish the current boundary planes (flag-ordered; the next step waits for
        the neighbours' copies on the device)."""
        self.state.p2p_refresh()

    def step(self, n_steps: int = 1) -> None:
        self.state.p2p_step(n_steps)
