"""Multi-rank parity check, launched with torch.distributed.run (one process per GPU, NCCL):
every rank steps its z-slab with halo exchange; rank 0 also steps the whole lattice on its own GPU and the
gathered slabs must be BIT-IDENTICAL to it (same kernels, exact halos).  Prints 'SLAB PARITY OK'."""
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def main():
    import torch
    import torch.distributed as dist

    import fdtd_b200
    from bench import build_workload
    from fdtd_b200.distributed import CudaSlabEngine, SlabStepper, partition_z

    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    lattice = tuple(int(v) for v in (sys.argv[1] if len(sys.argv) > 1 else "64,48,40").split(","))
    steps = int(sys.argv[2]) if len(sys.argv) > 2 else 25
    overlap = (sys.argv[3] != "0") if len(sys.argv) > 3 else True
    fused = (sys.argv[4] != "0") if len(sys.argv) > 4 else False
    transport = sys.argv[5] if len(sys.argv) > 5 else "nccl"     # nccl | p2p (NVLink peer memory, fused handles only)
    pml = int(sys.argv[6]) if len(sys.argv) > 6 else 0            # CPML cells per face (two-kernel handles over NCCL, or fused handles over p2p: DESIGN §4b)
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    zb, ze = partition_z(lattice[2], world)[rank]
    w = build_workload(lattice, zb, ze)
    # move the source next to a slab boundary so forcing, halos and probes interact
    cut = partition_z(lattice[2], world)[0][1]
    src_pt = (lattice[0] // 2, lattice[1] // 2, cut - 1)
    w.sources = [(src_pt, w.sources[0][1])]
    w.probes = [(lattice[0] // 2 + 3, lattice[1] // 2, cut), (lattice[0] // 2, lattice[1] // 2 + 2, cut - 1), (1, 1, lattice[2] - 1)]
    tab = w.source_table(steps)
    if pml:
        w.pml = fdtd_b200.PmlProfiles.graded(w.solver_config(), pml)
    be = fdtd_b200.FdtdB200Backend(device=local, z_slab=(zb, ze), fused_step=fused)
    st = be.create_instance(w.solver_config(), w.domain()).create_state()
    st.register_probes(w.probes)
    st.set_source_table(w.source_points(), tab)
    if transport == "p2p":
        from fdtd_b200.distributed import P2PSlabStepper

        stepper = P2PSlabStepper(st, rank, world)
        stepper.exchanges = steps
    else:
        stepper = SlabStepper(CudaSlabEngine(st), rank, world, overlap=overlap)
    stepper.step(steps)
    st.sync()
    torch.cuda.synchronize()
    e = torch.from_numpy(st.read_field(fdtd_b200.FieldComponent.E)).cuda()
    h = torch.from_numpy(st.read_field(fdtd_b200.FieldComponent.H)).cuda()
    pr = torch.from_numpy(st.read_probes(0, steps)).cuda()
    dist.all_reduce(pr)  # probes outside a slab read as zeros → the sum is the owner's sample
    sizes = [b - a for a, b in partition_z(lattice[2], world)]
    es = [torch.empty((n, lattice[1], lattice[0], 3), dtype=torch.float32, device="cuda") for n in sizes]
    hs = [torch.empty((n, lattice[1], lattice[0], 3), dtype=torch.float32, device="cuda") for n in sizes]
    dist.all_gather(es, e)
    dist.all_gather(hs, h)
    ok = True
    if rank == 0:
        wf = build_workload(lattice, 0, lattice[2])
        wf.sources, wf.probes, wf.pml = w.sources, w.probes, w.pml
        full = fdtd_b200.FdtdB200Backend(device=local).create_instance(wf.solver_config(), wf.domain()).create_state()
        full.register_probes(wf.probes)
        full.set_source_table(wf.source_points(), tab)
        full.step(steps)
        ef, hf = full.read_field(fdtd_b200.FieldComponent.E), full.read_field(fdtd_b200.FieldComponent.H)
        eg, hg = torch.cat(es).cpu().numpy(), torch.cat(hs).cpu().numpy()
        ok = np.array_equal(eg, ef) and np.array_equal(hg, hf) and np.array_equal(pr.cpu().numpy(), full.read_probes(0, steps))
        ok = ok and bool(np.any(ef)) and st.tick() == steps
        print(f"world={world} lattice={lattice} steps={steps} overlap={overlap} fused={st.uses_fused} transport={transport} pml={pml} exchanges={stepper.exchanges}", "SLAB PARITY OK" if ok else "SLAB PARITY FAILED", flush=True)
    dist.barrier()
    dist.destroy_process_group()
    sys.exit(0 if ok else 1)


if __name__ == "__main__":
    main()
