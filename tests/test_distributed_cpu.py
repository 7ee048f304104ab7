"""CPU (gloo, world_size 2 and 3) test of the z-slab halo-exchange orchestration in fdtd_b200/distributed.py.
The per-rank slab engine here is the numpy formulation of the reference update (tests/np_slab.py, test
infrastructure); the SlabStepper under test is the product's.  The decomposed run must equal the single-domain
oracle: same kernels + exact halos ⇒ identical up to f64 rounding."""
import os
import socket
import sys

import numpy as np
import pytest

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, size, steps, overlap, out_dir):
    sys.path.insert(0, ROOT)
    sys.path.insert(0, HERE)
    import torch.distributed as dist

    import oracle
    from fdtd_b200.distributed import SlabStepper, partition_z
    from np_slab import NumpySlab

    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    nx, ny, nz = size
    rng = np.random.default_rng(7)
    pal = np.array([[1, 0, 1, 0], [1.0, 0.0, 3.9, 0.0], [2.0, 0.5, 1.0, 0.0], [1.0, 0.0, 1.0, 10.0]])
    ids = rng.integers(0, 4, nx * ny * nz)
    dx, dt = (1.0, 0.5, 2.0), 0.2
    coef = np.array([oracle.update_coefficients(dt, 1.0, 1.0, pal[k]) for k in ids]).reshape(nz, ny, nx, 4)
    e0 = rng.standard_normal((nz, ny, nx, 3))
    h0 = rng.standard_normal((nz, ny, nx, 3))
    zb, ze = partition_z(nz, world)[rank]
    slab = NumpySlab(size, dx, dt, coef, zb, ze)
    for k, c in enumerate(("ex", "ey", "ez")):
        slab.f[c][1:-1] = e0[zb:ze, :, :, k]
    for k, c in enumerate(("hx", "hy", "hz")):
        slab.f[c][1:-1] = h0[zb:ze, :, :, k]
    stepper = SlabStepper(slab, rank, world, overlap=overlap)
    stepper.refresh_halos()
    src = (nx // 2, ny // 2, nz // 2)
    for s in range(steps):
        slab.set_forcing(src, (0.1 * s, -1.0, 2.0), (0.5, 0.0, -0.25 * s))
        stepper.step(1)
    np.savez(os.path.join(out_dir, f"rank{rank}.npz"), E=slab.field("E"), H=slab.field("H"), zb=zb, ze=ze,
             exchanges=stepper.exchanges, tick=slab.tick)
    dist.barrier()
    dist.destroy_process_group()


@pytest.mark.parametrize("world,overlap,size", [(2, True, (6, 5, 8)), (2, False, (6, 5, 8)), (3, True, (5, 4, 7)), (2, True, (4, 3, 2))])
def test_slab_stepper_matches_single_domain(tmp_path, oracle, world, overlap, size):
    import torch.multiprocessing as mp

    steps = 6
    port = _free_port()
    mp.spawn(_worker, args=(world, port, size, steps, overlap, str(tmp_path)), nprocs=world, join=True)

    nx, ny, nz = size
    rng = np.random.default_rng(7)
    pal = np.array([[1, 0, 1, 0], [1.0, 0.0, 3.9, 0.0], [2.0, 0.5, 1.0, 0.0], [1.0, 0.0, 1.0, 10.0]])
    ids = rng.integers(0, 4, nx * ny * nz)
    dx, dt = (1.0, 0.5, 2.0), 0.2
    ref = oracle.RefSolver(size, dx, dt)
    ref.set_materials(pal, ids.astype(np.uint32), 1.0, 1.0)
    ref.write_field(oracle.E, rng.standard_normal((nz, ny, nx, 3)))
    ref.write_field(oracle.H, rng.standard_normal((nz, ny, nx, 3)))
    src = (nx // 2, ny // 2, nz // 2)
    for s in range(steps):
        ref.begin_update()
        ref.set_forcing(src, (0.1 * s, -1.0, 2.0), (0.5, 0.0, -0.25 * s))
        ref.finish()
    parts = [np.load(os.path.join(str(tmp_path), f"rank{r}.npz")) for r in range(world)]
    assert [int(p["zb"]) for p in parts] == [a for a, _ in __import__("fdtd_b200.distributed", fromlist=["x"]).partition_z(nz, world)]
    e = np.concatenate([p["E"] for p in parts], axis=0)
    h = np.concatenate([p["H"] for p in parts], axis=0)
    assert np.allclose(e, ref.field(oracle.E), rtol=1e-13, atol=1e-13)
    assert np.allclose(h, ref.field(oracle.H), rtol=1e-13, atol=1e-13)
    # two neighbour exchanges per step (+2 for refresh_halos) on every rank that has a neighbour
    assert all(int(p["exchanges"]) == 2 * steps + 2 for p in parts) and all(int(p["tick"]) == steps for p in parts)


def test_partition_z():
    from fdtd_b200.distributed import partition_z

    assert partition_z(512, 1) == [(0, 512)]
    assert partition_z(1024, 8) == [(128 * r, 128 * (r + 1)) for r in range(8)]
    assert partition_z(10, 3) == [(0, 4), (4, 7), (7, 10)]
    with pytest.raises(ValueError):
        partition_z(2, 3)
