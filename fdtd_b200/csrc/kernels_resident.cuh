// Persistent multi-step kernel for lattices whose whole state is resident on chip (the lattices the reference application
// really runs: 100×100×1, cem-app/src/composer/mod.rs:841-869; 64³, BASELINE configs[0]).
//
// On such lattices a launch per half-step (or per step) is all overhead: the HBM-streaming kernels pay their TMA ring
// prologue for a march of one or a few planes (100×100×1: 8–14 µs per step), and nothing amortises the launch.  Here ONE
// launch runs n leapfrog steps (cem-solver/src/fdtd/cpu/mod.rs:416-515 × n):
//   * every thread owns CPT ∈ {1, 2, 4} fixed cells for the whole run and keeps their six field values, their four update
//     coefficients, their boundary flags and their forcing / probe slots IN REGISTERS — per half-step it loads only the six
//     neighbour values the curl needs (all independent: one L2 round trip), computes, and stores its three new values for
//     the neighbours to read;
//   * the H half-step and the E half-step are separated by a barrier across all CTAs: the hardware cluster barrier when the
//     lattice fits one thread-block cluster (≤ 16 CTAs of 512 threads × CPT cells), else a cooperative-groups grid barrier
//     over a cooperative launch (≤ one CTA per SM);
//   * neighbour values are read with ld.global.cg (from L2 — another SM wrote them), written with st.global.cg; the
//     fields never leave L2 and are updated in place, so the kernel mixes freely with every other entry point;
//   * forcing rows are consumed one per step from the device-resident table, probe rows go into the ring exactly as the
//     other kernels write them — `fdtd_b200_step(h, n)` is one launch.
//
// Arithmetic: the same helpers as every other kernel generation (deriv / leap, kernels.cuh) in the same order, so exact mode is
// bit-identical to the fp32 oracle and fast mode to the other fast-mode kernels (the Neumann select after the product keeps
// nvcc from contracting it into the curl's subtraction, as in k_update_*_vec4).
#pragma once

#include <cooperative_groups.h>

#include "kernels.cuh"
#include "kernels_bulk.cuh"   // smem_u32

namespace fdtd_b200 {

constexpr int kResidentThreads = 512;   // one CTA per SM, up to 128 registers per thread for the resident cell state

struct ResidentView {
    LatticeView L;
    int n_steps;
    // forcing table rows [0, src_rows) are consumed by steps 0.. of this launch; later steps run unforced
    int n_src;
    const long long* src_cells;
    const float* src_vals;     // [src_rows][n_src][6]
    int src_rows;
    // probes: row of step s = ring + ((tick0 + s) % ring_steps) * n_probes_total * 6
    int n_probes;
    const long long* probe_cells;
    const int* probe_slot;
    float* ring;
    unsigned long long tick0;
    unsigned int ring_steps;
    unsigned int n_probes_total;
    // grid barrier: one 8-byte slot per CTA holding the last epoch it reached; this launch uses epochs epoch0 + 1 …
    unsigned long long* sync_flags;
    unsigned long long epoch0;
    int barrier_mode;          // grid regime: 0 cooperative-groups grid.sync(), 1 all-to-all flags (measured slower: the pollers' hot lines)
    long long* dbg;            // FDTD_B200_RESIDENT_DEBUG: clock64 stamps of thread 0 of every CTA, [cta][step < 8][8]
};

// Barrier across all CTAs of the launch.
//  CLUSTER: the hardware cluster barrier; release/acquire at cluster scope orders the global stores of every CTA of the
//           cluster before the loads that follow.
//  grid:    an all-to-all flag barrier over a cooperative launch (all CTAs co-resident).  Every CTA publishes the epoch it has
//           reached in its own 8-byte slot (one fence + one store by thread 0 after the CTA's stores were issued) and polls
//           everybody's slot, one slot per thread.  Latency = one store + one load through L2, whatever the CTA count;
//           cooperative_groups' grid.sync() funnels every CTA through an atomic on ONE address, which the L2 serialises —
//           measured 3400 cycles per barrier with 128 CTAs (profiles/r02_ncu_resident.md), half of the step.
template <bool CLUSTER>
__device__ __forceinline__ void resident_barrier(unsigned long long* flags, unsigned long long epoch, int mode) {
    if (CLUSTER) {
        asm volatile("barrier.cluster.arrive.release.aligned;\n" ::: "memory");
        asm volatile("barrier.cluster.wait.acquire.aligned;\n" ::: "memory");
    } else if (mode == 0) {
        cooperative_groups::this_grid().sync();
    } else {
        __syncthreads();
        if (threadIdx.x == 0) {
            __threadfence();
            asm volatile("st.relaxed.gpu.global.u64 [%0], %1;\n" :: "l"(flags + blockIdx.x), "l"(epoch) : "memory");
        }
        if (threadIdx.x < gridDim.x) {
            unsigned long long seen;
            do {
                asm volatile("ld.acquire.gpu.global.u64 %0, [%1];\n" : "=l"(seen) : "l"(flags + threadIdx.x) : "memory");
            } while (seen < epoch);
        }
        __syncthreads();
    }
}

template <int EXACT>
__device__ __forceinline__ float deriv_sel(bool in, float hi, float lo, float dx, float inv_dx) {
    return in ? deriv<EXACT>(hi, lo, dx, inv_dx) : 0.f;
}

template <int EXACT, typename IdxT, bool CLUSTER, int CPT>
__global__ void __launch_bounds__(kResidentThreads, 1)
k_steps_resident(const ResidentView V) {
    const LatticeView& L = V.L;
    const int nx = L.nx, ny = L.ny, nz = L.nz, nxl = L.nxl;
    const int sy = nx, sz = nx * ny;
    const int n = sz * nz;                                   // resident lattices are far below 2^31 cells
    const IdxT* __restrict__ mat = static_cast<const IdxT*>(L.mat);

    // ---- per-cell state that never changes during the launch, and the cell's own fields: registers
    enum : unsigned { LIVE = 1u, XM = 2u, YM = 4u, ZM = 8u, XP = 16u, YP = 32u, ZP = 64u };
    int idx[CPT];
    unsigned flags[CPT];
    float4 coef[CPT];
    float ex[CPT], ey[CPT], ez[CPT], hx[CPT], hy[CPT], hz[CPT];
    int src_k[CPT];        // this cell's column in the forcing table, −1: none
    int probe_k[CPT];      // first entry of the (sorted) probe list that samples this cell, −1: none
#pragma unroll
    for (int k = 0; k < CPT; ++k) {
        const int i = (blockIdx.x * CPT + k) * kResidentThreads + threadIdx.x;   // consecutive threads, consecutive cells
        idx[k] = i;
        flags[k] = 0; src_k[k] = -1; probe_k[k] = -1;
        coef[k] = make_float4(0.f, 0.f, 0.f, 0.f);
        ex[k] = ey[k] = ez[k] = hx[k] = hy[k] = hz[k] = 0.f;
        if (i < n) {
            const int z = i / sz;
            const int r = i - z * sz;
            const int y = r / nx;
            const int x = r - y * nx;
            if (x < nxl) {                                   // pad cells of a padded row stay 0 and are never touched
                flags[k] = LIVE | (x > 0 ? XM : 0u) | (y > 0 ? YM : 0u) | (z > 0 ? ZM : 0u) |
                           (x + 1 < nxl ? XP : 0u) | (y + 1 < ny ? YP : 0u) | (z + 1 < nz ? ZP : 0u);
                coef[k] = __ldg(L.table + mat[i]);
                ex[k] = L.ex[i]; ey[k] = L.ey[i]; ez[k] = L.ez[i];
                hx[k] = L.hx[i]; hy[k] = L.hy[i]; hz[k] = L.hz[i];
                if (V.n_src) {
                    const int s = lower_bound_ll(V.src_cells, V.n_src, static_cast<long long>(i));
                    if (s < V.n_src && __ldg(V.src_cells + s) == static_cast<long long>(i)) src_k[k] = s;
                }
                if (V.n_probes) {
                    const int s = lower_bound_ll(V.probe_cells, V.n_probes, static_cast<long long>(i));
                    if (s < V.n_probes && __ldg(V.probe_cells + s) == static_cast<long long>(i)) probe_k[k] = s;
                }
            }
        }
    }

    for (int s = 0; s < V.n_steps; ++s) {
        const bool forced = s < V.src_rows;
        const float* __restrict__ vals = V.src_vals + static_cast<size_t>(s) * V.n_src * 6;

        if (V.dbg && threadIdx.x == 0 && s < 8) V.dbg[(blockIdx.x * 8 + s) * 8 + 0] = clock64();
        // ---- H' = d_a·H + d_b·(−∇×E − M): backward differences, zero where x_i == 0 (update.wgsl:56-91, 155-165)
        float n0[CPT], n1[CPT], n2[CPT], n3[CPT], n4[CPT], n5[CPT], f0[CPT], f1[CPT], f2[CPT];
#pragma unroll
        for (int k = 0; k < CPT; ++k) {                      // every load of the half-step is issued before the first use
            const int i = idx[k];
            const unsigned f = flags[k];
            n0[k] = (f & YM) ? __ldcg(L.ez + i - sy) : 0.f; n1[k] = (f & YM) ? __ldcg(L.ex + i - sy) : 0.f;
            n2[k] = (f & ZM) ? __ldcg(L.ey + i - sz) : 0.f; n3[k] = (f & ZM) ? __ldcg(L.ex + i - sz) : 0.f;
            n4[k] = (f & XM) ? __ldcg(L.ez + i - 1) : 0.f;  n5[k] = (f & XM) ? __ldcg(L.ey + i - 1) : 0.f;
            const bool src = forced && src_k[k] >= 0;
            f0[k] = src ? vals[6 * src_k[k] + 3] : 0.f; f1[k] = src ? vals[6 * src_k[k] + 4] : 0.f; f2[k] = src ? vals[6 * src_k[k] + 5] : 0.f;
        }
#pragma unroll
        for (int k = 0; k < CPT; ++k) {
            const unsigned f = flags[k];
            const float dez_dy = deriv_sel<EXACT>(f & YM, ez[k], n0[k], L.dx[1], L.inv_dx[1]);
            const float dex_dy = deriv_sel<EXACT>(f & YM, ex[k], n1[k], L.dx[1], L.inv_dx[1]);
            const float dey_dz = deriv_sel<EXACT>(f & ZM, ey[k], n2[k], L.dx[2], L.inv_dx[2]);
            const float dex_dz = deriv_sel<EXACT>(f & ZM, ex[k], n3[k], L.dx[2], L.inv_dx[2]);
            const float dez_dx = deriv_sel<EXACT>(f & XM, ez[k], n4[k], L.dx[0], L.inv_dx[0]);
            const float dey_dx = deriv_sel<EXACT>(f & XM, ey[k], n5[k], L.dx[0], L.inv_dx[0]);
            const float cx = dez_dy - dey_dz;
            const float cy = dex_dz - dez_dx;
            const float cz = dey_dx - dex_dy;
            if (f & LIVE) {
                hx[k] = leap<EXACT>(coef[k].z, hx[k], coef[k].w, -cx, f0[k]);
                hy[k] = leap<EXACT>(coef[k].z, hy[k], coef[k].w, -cy, f1[k]);
                hz[k] = leap<EXACT>(coef[k].z, hz[k], coef[k].w, -cz, f2[k]);
                __stcg(L.hx + idx[k], hx[k]); __stcg(L.hy + idx[k], hy[k]); __stcg(L.hz + idx[k], hz[k]);
            }
        }
        if (V.dbg && threadIdx.x == 0 && s < 8) V.dbg[(blockIdx.x * 8 + s) * 8 + 2] = clock64();
        resident_barrier<CLUSTER>(V.sync_flags, V.epoch0 + 2 * s + 1, V.barrier_mode);
        if (V.dbg && threadIdx.x == 0 && s < 8) V.dbg[(blockIdx.x * 8 + s) * 8 + 3] = clock64();

        // ---- E' = c_a·E + c_b·(∇×H' − J): forward differences on the new H, zero where x_i + 1 == n_i (update.wgsl:94-129, 167-177)
#pragma unroll
        for (int k = 0; k < CPT; ++k) {
            const int i = idx[k];
            const unsigned f = flags[k];
            n0[k] = (f & YP) ? __ldcg(L.hz + i + sy) : 0.f; n1[k] = (f & YP) ? __ldcg(L.hx + i + sy) : 0.f;
            n2[k] = (f & ZP) ? __ldcg(L.hy + i + sz) : 0.f; n3[k] = (f & ZP) ? __ldcg(L.hx + i + sz) : 0.f;
            n4[k] = (f & XP) ? __ldcg(L.hz + i + 1) : 0.f;  n5[k] = (f & XP) ? __ldcg(L.hy + i + 1) : 0.f;
            const bool src = forced && src_k[k] >= 0;
            f0[k] = src ? vals[6 * src_k[k]] : 0.f; f1[k] = src ? vals[6 * src_k[k] + 1] : 0.f; f2[k] = src ? vals[6 * src_k[k] + 2] : 0.f;
        }
        float* __restrict__ row = V.n_probes ? V.ring + ((V.tick0 + s) % V.ring_steps) * static_cast<size_t>(V.n_probes_total) * 6 : nullptr;
#pragma unroll
        for (int k = 0; k < CPT; ++k) {
            const unsigned f = flags[k];
            const float dhz_dy = deriv_sel<EXACT>(f & YP, n0[k], hz[k], L.dx[1], L.inv_dx[1]);
            const float dhx_dy = deriv_sel<EXACT>(f & YP, n1[k], hx[k], L.dx[1], L.inv_dx[1]);
            const float dhy_dz = deriv_sel<EXACT>(f & ZP, n2[k], hy[k], L.dx[2], L.inv_dx[2]);
            const float dhx_dz = deriv_sel<EXACT>(f & ZP, n3[k], hx[k], L.dx[2], L.inv_dx[2]);
            const float dhz_dx = deriv_sel<EXACT>(f & XP, n4[k], hz[k], L.dx[0], L.inv_dx[0]);
            const float dhy_dx = deriv_sel<EXACT>(f & XP, n5[k], hy[k], L.dx[0], L.inv_dx[0]);
            const float cx = dhz_dy - dhy_dz;
            const float cy = dhx_dz - dhz_dx;
            const float cz = dhy_dx - dhx_dy;
            if (f & LIVE) {
                ex[k] = leap<EXACT>(coef[k].x, ex[k], coef[k].y, cx, f0[k]);
                ey[k] = leap<EXACT>(coef[k].x, ey[k], coef[k].y, cy, f1[k]);
                ez[k] = leap<EXACT>(coef[k].x, ez[k], coef[k].y, cz, f2[k]);
                __stcg(L.ex + idx[k], ex[k]); __stcg(L.ey + idx[k], ey[k]); __stcg(L.ez + idx[k], ez[k]);
                if (probe_k[k] >= 0) {
                    for (int p = probe_k[k]; p < V.n_probes && __ldg(V.probe_cells + p) == static_cast<long long>(idx[k]); ++p) {
                        float* o = row + 6 * static_cast<size_t>(__ldg(V.probe_slot + p));
                        o[0] = ex[k]; o[1] = ey[k]; o[2] = ez[k]; o[3] = hx[k]; o[4] = hy[k]; o[5] = hz[k];
                    }
                }
            }
        }
        if (V.dbg && threadIdx.x == 0 && s < 8) V.dbg[(blockIdx.x * 8 + s) * 8 + 6] = clock64();
        if (s + 1 < V.n_steps) resident_barrier<CLUSTER>(V.sync_flags, V.epoch0 + 2 * s + 2, V.barrier_mode);
        if (V.dbg && threadIdx.x == 0 && s < 8) V.dbg[(blockIdx.x * 8 + s) * 8 + 7] = clock64();
    }
}


// ---- the smallest lattices: one thread-block cluster, fields exchanged through DISTRIBUTED SHARED MEMORY ---------------------------
// With global memory as the exchange medium a barrier costs ≈ 2000 cycles even inside a cluster: its release half has to wait
// until the CTA's field stores are acknowledged by L2 (MEMBAR.ALL.GPU, the top stall of the first version,
// profiles/r02_ncu_resident.md).  Here the six fields of the CTA's cells live in its shared memory for the whole launch; a
// neighbour cell owned by another CTA of the cluster is read from that CTA's shared memory (ld.shared::cluster, ≈ 215 cycles),
// nothing is stored to global memory between two barriers, and `barrier.cluster` (≈ 380 cycles) is all that separates the half-
// steps.  Probe rows are staged in shared memory too and flushed, like the fields, when the launch ends (the host bounds the
// steps per launch by the staging capacity: ResidentView::n_steps ≤ kProbeStageFloats / (6 · probes of the busiest CTA)).
constexpr int kProbeStageFloats = 4096;   // 16 KB next to the ≤ 24 KB of fields

__device__ __forceinline__ unsigned map_to_cta(unsigned local_addr, unsigned rank) {
    unsigned out;
    asm volatile("mapa.shared::cluster.u32 %0, %1, %2;\n" : "=r"(out) : "r"(local_addr), "r"(rank));
    return out;
}
// Barrier of the distributed-shared-memory kernel.  `barrier.cluster.arrive.release` compiles to MEMBAR.ALL.GPU + the arrive and
// costs ≈ 1400 cycles even with no global store in flight (measured with the clock64 stamps below: 880-cycle half-steps between
// 1400-cycle barriers).  The kernel exchanges data through shared memory only, so the release it needs is narrower: the CTA's
// own st.shared must have been performed before any peer reads them.  BAR.SYNC drains the CTA's pending shared-memory stores
// (they are then in this SM's shared memory, which is where peers read them from), after which a RELAXED cluster arrive/wait
// pair is enough; the reverse hazard (peers must have read my old values before I overwrite them) is covered by data
// dependence — a peer's stores, and therefore its arrive, come after the loads whose values they are computed from.
// FDTD_B200_RESIDENT_BARRIER=2 selects the release/acquire form for comparison; the parity tests run both.
__device__ __forceinline__ void dsmem_barrier(int mode) {
    if (mode == 2) {
        asm volatile("barrier.cluster.arrive.release.aligned;\n" ::: "memory");
        asm volatile("barrier.cluster.wait.acquire.aligned;\n" ::: "memory");
    } else {
        __syncthreads();
        asm volatile("barrier.cluster.arrive.relaxed.aligned;\n" ::: "memory");
        asm volatile("barrier.cluster.wait.aligned;\n" ::: "memory");
    }
}

__device__ __forceinline__ float ld_dsmem(unsigned addr) {
    float v;
    asm volatile("ld.shared::cluster.f32 %0, [%1];\n" : "=f"(v) : "r"(addr) : "memory");
    return v;
}

template <int EXACT, typename IdxT, int CPT>
__global__ void __launch_bounds__(kResidentThreads, 1)
k_steps_cluster(const ResidentView V) {
    constexpr int CELLS = kResidentThreads * CPT;            // cells of one CTA: [rank·CELLS, (rank+1)·CELLS)
    __shared__ float s_f[6][CELLS];                          // ex ey ez hx hy hz
    __shared__ float s_probe[kProbeStageFloats];             // [step][probe entry of this CTA][6]
    const LatticeView& L = V.L;
    const int nx = L.nx, ny = L.ny, nz = L.nz, nxl = L.nxl;
    const int sy = nx, sz = nx * ny;
    const int n = sz * nz;
    const IdxT* __restrict__ mat = static_cast<const IdxT*>(L.mat);
    const int cta_first = blockIdx.x * CELLS;                // the grid is one cluster: blockIdx.x is the rank in the cluster

    enum : unsigned { LIVE = 1u, XM = 2u, YM = 4u, ZM = 8u, XP = 16u, YP = 32u, ZP = 64u };
    int idx[CPT];
    unsigned flags[CPT];
    float4 coef[CPT];
    float ex[CPT], ey[CPT], ez[CPT], hx[CPT], hy[CPT], hz[CPT];
    int src_k[CPT], probe_k[CPT];
    unsigned a_xm[CPT], a_ym[CPT], a_zm[CPT], a_xp[CPT], a_yp[CPT], a_zp[CPT];   // shared::cluster address of the neighbour's ex slot
    // the probe entries of this CTA's cells are one contiguous run [p_first, p_first + p_count) of the sorted list
    const int p_first = V.n_probes ? lower_bound_ll(V.probe_cells, V.n_probes, static_cast<long long>(cta_first)) : 0;
    const int p_count = V.n_probes ? lower_bound_ll(V.probe_cells, V.n_probes, static_cast<long long>(cta_first) + CELLS) - p_first : 0;

    auto neighbour = [&](int j) -> unsigned {                // slot of cell j's ex in its owner's shared memory
        const unsigned r = static_cast<unsigned>(j / CELLS);
        return map_to_cta(smem_u32(&s_f[0][j - static_cast<int>(r) * CELLS]), r);
    };
#pragma unroll
    for (int k = 0; k < CPT; ++k) {
        const int o = k * kResidentThreads + threadIdx.x;
        const int i = cta_first + o;
        idx[k] = i;
        flags[k] = 0; src_k[k] = -1; probe_k[k] = -1;
        coef[k] = make_float4(0.f, 0.f, 0.f, 0.f);
        ex[k] = ey[k] = ez[k] = hx[k] = hy[k] = hz[k] = 0.f;
        a_xm[k] = a_ym[k] = a_zm[k] = a_xp[k] = a_yp[k] = a_zp[k] = smem_u32(&s_f[0][o]);
        if (i < n) {
            const int z = i / sz;
            const int r = i - z * sz;
            const int y = r / nx;
            const int x = r - y * nx;
            if (x < nxl) {
                flags[k] = LIVE | (x > 0 ? XM : 0u) | (y > 0 ? YM : 0u) | (z > 0 ? ZM : 0u) |
                           (x + 1 < nxl ? XP : 0u) | (y + 1 < ny ? YP : 0u) | (z + 1 < nz ? ZP : 0u);
                coef[k] = __ldg(L.table + mat[i]);
                ex[k] = L.ex[i]; ey[k] = L.ey[i]; ez[k] = L.ez[i];
                hx[k] = L.hx[i]; hy[k] = L.hy[i]; hz[k] = L.hz[i];
                if (flags[k] & XM) a_xm[k] = neighbour(i - 1);
                if (flags[k] & YM) a_ym[k] = neighbour(i - sy);
                if (flags[k] & ZM) a_zm[k] = neighbour(i - sz);
                if (flags[k] & XP) a_xp[k] = neighbour(i + 1);
                if (flags[k] & YP) a_yp[k] = neighbour(i + sy);
                if (flags[k] & ZP) a_zp[k] = neighbour(i + sz);
                if (V.n_src) {
                    const int q = lower_bound_ll(V.src_cells, V.n_src, static_cast<long long>(i));
                    if (q < V.n_src && __ldg(V.src_cells + q) == static_cast<long long>(i)) src_k[k] = q;
                }
                if (p_count) {
                    const int q = lower_bound_ll(V.probe_cells, V.n_probes, static_cast<long long>(i));
                    if (q < V.n_probes && __ldg(V.probe_cells + q) == static_cast<long long>(i)) probe_k[k] = q;
                }
            }
        }
        s_f[0][o] = ex[k]; s_f[1][o] = ey[k]; s_f[2][o] = ez[k];
        s_f[3][o] = hx[k]; s_f[4][o] = hy[k]; s_f[5][o] = hz[k];
    }
    constexpr unsigned COMP = CELLS * sizeof(float);         // byte distance between two components of one cell
    dsmem_barrier(V.barrier_mode);                   // every CTA's shared memory holds its cells

    for (int s = 0; s < V.n_steps; ++s) {
        const bool forced = s < V.src_rows;
        const float* __restrict__ vals = V.src_vals + static_cast<size_t>(s) * V.n_src * 6;
        if (V.dbg && threadIdx.x == 0 && s < 8) V.dbg[(blockIdx.x * 8 + s) * 8 + 0] = clock64();

        // ---- H half-step (update.wgsl:56-91, 155-165): reads E of the lower neighbours
        float n0[CPT], n1[CPT], n2[CPT], n3[CPT], n4[CPT], n5[CPT], f0[CPT], f1[CPT], f2[CPT];
#pragma unroll
        for (int k = 0; k < CPT; ++k) {
            const unsigned f = flags[k];
            n0[k] = (f & YM) ? ld_dsmem(a_ym[k] + 2 * COMP) : 0.f; n1[k] = (f & YM) ? ld_dsmem(a_ym[k]) : 0.f;              // ez, ex at y−1
            n2[k] = (f & ZM) ? ld_dsmem(a_zm[k] + COMP) : 0.f;     n3[k] = (f & ZM) ? ld_dsmem(a_zm[k]) : 0.f;              // ey, ex at z−1
            n4[k] = (f & XM) ? ld_dsmem(a_xm[k] + 2 * COMP) : 0.f; n5[k] = (f & XM) ? ld_dsmem(a_xm[k] + COMP) : 0.f;       // ez, ey at x−1
            const bool src = forced && src_k[k] >= 0;
            f0[k] = src ? vals[6 * src_k[k] + 3] : 0.f; f1[k] = src ? vals[6 * src_k[k] + 4] : 0.f; f2[k] = src ? vals[6 * src_k[k] + 5] : 0.f;
        }
#pragma unroll
        for (int k = 0; k < CPT; ++k) {
            const unsigned f = flags[k];
            const float dez_dy = deriv_sel<EXACT>(f & YM, ez[k], n0[k], L.dx[1], L.inv_dx[1]);
            const float dex_dy = deriv_sel<EXACT>(f & YM, ex[k], n1[k], L.dx[1], L.inv_dx[1]);
            const float dey_dz = deriv_sel<EXACT>(f & ZM, ey[k], n2[k], L.dx[2], L.inv_dx[2]);
            const float dex_dz = deriv_sel<EXACT>(f & ZM, ex[k], n3[k], L.dx[2], L.inv_dx[2]);
            const float dez_dx = deriv_sel<EXACT>(f & XM, ez[k], n4[k], L.dx[0], L.inv_dx[0]);
            const float dey_dx = deriv_sel<EXACT>(f & XM, ey[k], n5[k], L.dx[0], L.inv_dx[0]);
            const float cx = dez_dy - dey_dz;
            const float cy = dex_dz - dez_dx;
            const float cz = dey_dx - dex_dy;
            if (f & LIVE) {
                const int o = k * kResidentThreads + threadIdx.x;
                hx[k] = leap<EXACT>(coef[k].z, hx[k], coef[k].w, -cx, f0[k]);
                hy[k] = leap<EXACT>(coef[k].z, hy[k], coef[k].w, -cy, f1[k]);
                hz[k] = leap<EXACT>(coef[k].z, hz[k], coef[k].w, -cz, f2[k]);
                s_f[3][o] = hx[k]; s_f[4][o] = hy[k]; s_f[5][o] = hz[k];
            }
        }
        if (V.dbg && threadIdx.x == 0 && s < 8) V.dbg[(blockIdx.x * 8 + s) * 8 + 2] = clock64();
        dsmem_barrier(V.barrier_mode);
        if (V.dbg && threadIdx.x == 0 && s < 8) V.dbg[(blockIdx.x * 8 + s) * 8 + 3] = clock64();

        // ---- E half-step (update.wgsl:94-129, 167-177): reads the new H of the upper neighbours
#pragma unroll
        for (int k = 0; k < CPT; ++k) {
            const unsigned f = flags[k];
            n0[k] = (f & YP) ? ld_dsmem(a_yp[k] + 5 * COMP) : 0.f; n1[k] = (f & YP) ? ld_dsmem(a_yp[k] + 3 * COMP) : 0.f;   // hz, hx at y+1
            n2[k] = (f & ZP) ? ld_dsmem(a_zp[k] + 4 * COMP) : 0.f; n3[k] = (f & ZP) ? ld_dsmem(a_zp[k] + 3 * COMP) : 0.f;   // hy, hx at z+1
            n4[k] = (f & XP) ? ld_dsmem(a_xp[k] + 5 * COMP) : 0.f; n5[k] = (f & XP) ? ld_dsmem(a_xp[k] + 4 * COMP) : 0.f;   // hz, hy at x+1
            const bool src = forced && src_k[k] >= 0;
            f0[k] = src ? vals[6 * src_k[k]] : 0.f; f1[k] = src ? vals[6 * src_k[k] + 1] : 0.f; f2[k] = src ? vals[6 * src_k[k] + 2] : 0.f;
        }
#pragma unroll
        for (int k = 0; k < CPT; ++k) {
            const unsigned f = flags[k];
            const float dhz_dy = deriv_sel<EXACT>(f & YP, n0[k], hz[k], L.dx[1], L.inv_dx[1]);
            const float dhx_dy = deriv_sel<EXACT>(f & YP, n1[k], hx[k], L.dx[1], L.inv_dx[1]);
            const float dhy_dz = deriv_sel<EXACT>(f & ZP, n2[k], hy[k], L.dx[2], L.inv_dx[2]);
            const float dhx_dz = deriv_sel<EXACT>(f & ZP, n3[k], hx[k], L.dx[2], L.inv_dx[2]);
            const float dhz_dx = deriv_sel<EXACT>(f & XP, n4[k], hz[k], L.dx[0], L.inv_dx[0]);
            const float dhy_dx = deriv_sel<EXACT>(f & XP, n5[k], hy[k], L.dx[0], L.inv_dx[0]);
            const float cx = dhz_dy - dhy_dz;
            const float cy = dhx_dz - dhz_dx;
            const float cz = dhy_dx - dhx_dy;
            if (f & LIVE) {
                const int o = k * kResidentThreads + threadIdx.x;
                ex[k] = leap<EXACT>(coef[k].x, ex[k], coef[k].y, cx, f0[k]);
                ey[k] = leap<EXACT>(coef[k].x, ey[k], coef[k].y, cy, f1[k]);
                ez[k] = leap<EXACT>(coef[k].x, ez[k], coef[k].y, cz, f2[k]);
                s_f[0][o] = ex[k]; s_f[1][o] = ey[k]; s_f[2][o] = ez[k];
                if (probe_k[k] >= 0) {
                    for (int p = probe_k[k]; p < V.n_probes && __ldg(V.probe_cells + p) == static_cast<long long>(idx[k]); ++p) {
                        float* q = s_probe + (static_cast<size_t>(s) * p_count + (p - p_first)) * 6;
                        q[0] = ex[k]; q[1] = ey[k]; q[2] = ez[k]; q[3] = hx[k]; q[4] = hy[k]; q[5] = hz[k];
                    }
                }
            }
        }
        if (V.dbg && threadIdx.x == 0 && s < 8) V.dbg[(blockIdx.x * 8 + s) * 8 + 6] = clock64();
        dsmem_barrier(V.barrier_mode);               // also after the last step: no CTA may exit while its shared memory is being read
        if (V.dbg && threadIdx.x == 0 && s < 8) V.dbg[(blockIdx.x * 8 + s) * 8 + 7] = clock64();
    }

    // ---- the launch ends: fields and staged probe rows go back to global memory
#pragma unroll
    for (int k = 0; k < CPT; ++k) {
        if (flags[k] & LIVE) {
            const int i = idx[k];
            L.ex[i] = ex[k]; L.ey[i] = ey[k]; L.ez[i] = ez[k];
            L.hx[i] = hx[k]; L.hy[i] = hy[k]; L.hz[i] = hz[k];
        }
    }
    __syncthreads();
    for (int t = threadIdx.x; t < V.n_steps * p_count; t += kResidentThreads) {
        const int s = t / p_count, q = t - s * p_count;
        float* o = V.ring + ((V.tick0 + s) % V.ring_steps) * static_cast<size_t>(V.n_probes_total) * 6 + 6 * static_cast<size_t>(__ldg(V.probe_slot + p_first + q));
        const float* src = s_probe + static_cast<size_t>(t) * 6;
#pragma unroll
        for (int c = 0; c < 6; ++c) o[c] = src[c];
    }
}

}  // namespace fdtd_b200
