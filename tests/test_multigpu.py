"""Multi-GPU parity (needs >= 2 CUDA devices; skipped on a single-GPU box): z-slab ranks over NCCL must be
bit-identical to the single-domain run.  The same orchestration is covered on CPU by test_distributed_cpu.py."""
import os
import subprocess
import sys

import pytest

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.mark.parametrize("overlap,fused,transport", [("1", "0", "nccl"), ("0", "0", "nccl"), ("1", "1", "nccl"), ("0", "1", "nccl"), ("1", "1", "p2p")])
def test_two_ranks_bit_identical_to_one(b200, overlap, fused, transport):
    n = b200.lib().fdtd_b200_device_count()
    if n < 2:
        pytest.skip("needs two GPUs")
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=2", "--master-addr", "127.0.0.1",
           "--master-port", "29571", os.path.join(ROOT, "tools", "slab_parity.py"), "128,48,40" if fused == "1" else "64,48,40", "25", overlap, fused, transport]
    res = subprocess.run(cmd, capture_output=True, text=True, timeout=600)
    assert res.returncode == 0 and "SLAB PARITY OK" in res.stdout, res.stdout[-2000:] + res.stderr[-2000:]


def test_two_ranks_with_cpml_bit_identical_to_one(b200):
    """CPML slabs cross the cut (x and y layers) and sit on the outer z faces of both ranks; two-kernel handles over NCCL."""
    if b200.lib().fdtd_b200_device_count() < 2:
        pytest.skip("needs two GPUs")
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=2", "--master-addr", "127.0.0.1",
           "--master-port", "29572", os.path.join(ROOT, "tools", "slab_parity.py"), "64,48,40", "25", "1", "0", "nccl", "6"]
    res = subprocess.run(cmd, capture_output=True, text=True, timeout=600)
    assert res.returncode == 0 and "SLAB PARITY OK" in res.stdout and "pml=6" in res.stdout, res.stdout[-2000:] + res.stderr[-2000:]


def test_two_ranks_with_cpml_inside_the_fused_kernel_over_peer_memory(b200):
    """Fused z-slabs with the CPML advanced inside the step kernel; ψ_H of each slab's first plane travels to the slab below
    with the boundary planes (CUDA IPC peer memory).  Bit-identical to the single-domain two-kernel run with the slab passes."""
    if b200.lib().fdtd_b200_device_count() < 2:
        pytest.skip("needs two GPUs")
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=2", "--master-addr", "127.0.0.1",
           "--master-port", "29573", os.path.join(ROOT, "tools", "slab_parity.py"), "128,48,40", "25", "1", "1", "p2p", "6"]
    res = subprocess.run(cmd, capture_output=True, text=True, timeout=600)
    assert res.returncode == 0 and "SLAB PARITY OK" in res.stdout and "pml=6" in res.stdout and "fused=True" in res.stdout, res.stdout[-2000:] + res.stderr[-2000:]


@pytest.mark.parametrize("lattice,fused,transport,pml", [("200,48,40", "1", "p2p", "0"), ("200,48,40", "1", "nccl", "0"), ("90,48,40", "0", "nccl", "0"),
                                                       ("200,48,40", "1", "p2p", "6")])
def test_two_ranks_with_padded_rows_bit_identical_to_one(b200, lattice, fused, transport, pml):
    """Lattices whose nx is not a multiple of 16 (rows padded inside the library): halo planes, the pushed index
    plane and the pushed ψ planes carry the pitch on both ranks."""
    if b200.lib().fdtd_b200_device_count() < 2:
        pytest.skip("needs two GPUs")
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=2", "--master-addr", "127.0.0.1",
           "--master-port", "29574", os.path.join(ROOT, "tools", "slab_parity.py"), lattice, "25", "1", fused, transport, pml]
    res = subprocess.run(cmd, capture_output=True, text=True, timeout=600)
    assert res.returncode == 0 and "SLAB PARITY OK" in res.stdout, res.stdout[-2000:] + res.stderr[-2000:]
