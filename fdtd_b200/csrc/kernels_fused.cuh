// Fused leapfrog step (north_star item 4, "temporal blocking" of the two half-steps): ONE pass over the lattice
// computes H' = d_a·H + d_b·(−∇×E − M) and E' = c_a·E + c_b·(∇×H' − J) (cem-solver/src/fdtd/cpu/mod.rs:416-515,
// fdtd/wgpu/update.wgsl:56-177), so a cell update moves 6 fp32 reads + 6 fp32 writes + 1 B index = 49 B instead of 74 B.
//
// E'(p) needs H' at p, p+x, p+y, p+z, and a neighbouring CTA cannot see this CTA's fresh H' in time, so every CTA
// recomputes H' one row, one column and one plane beyond the cells whose E' it owns (overlapped tiling): a CTA of BY
// thread rows computes H' on BY rows and E' on the lower BY-1 of them (tiles advance by BY-1 rows; the top thread row
// has no E' work and is the TMA producer instead), and every row also computes the single H'y,H'z cell to its right.  That recomputation reads
// the OLD fields of cells another CTA is updating, hence the step is out of place: fields are ping-pong buffered
// (in → out) and the handle swaps the two sets after the launch.
//
// Data movement: old E (3 planes resident: z-1, z, z+1 in flight) and old H + material index (2 planes) are streamed
// into shared-memory rings by 3-D TMA tensor copies (cp.async.bulk.tensor.3d, one box per component and plane — SASS
// UTMALDG; out-of-lattice halo elements are zero-filled by the TMA unit) completing on mbarriers; the fresh H' of two consecutive
// planes is exchanged between threads through a double-buffered shared-memory tile.  One __syncthreads per plane.
// Arithmetic = the helpers of kernels.cuh → bit-identical to the two-kernel path.
#pragma once

#include <type_traits>

#include <cuda.h>  // CUtensorMap (types only; the encoder is fetched with cudaGetDriverEntryPoint)

#include "kernels_bulk.cuh"

namespace fdtd_b200 {

// Tensor maps of the step's input arrays (3-D: x, y, z; boxes of one plane of a CTA tile incl. halos).  The field maps
// start at the lower ghost plane (z index = plane + 1) so that slab handles can fetch their neighbours' planes.
struct FusedMaps {
    CUtensorMap e[3];   // old Ex, Ey, Ez : box (PE, E_ROWS, 1)
    CUtensorMap h[3];   // old Hx, Hy, Hz : box (PE, H_ROWS, 1)
    CUtensorMap mat;    // material index : box (PM, H_ROWS, 1)
};

__device__ __forceinline__ void tma_load_3d(void* dst_smem, const CUtensorMap* map, int x, int y, int z, uint64_t* bar) {
    asm volatile("cp.async.bulk.tensor.3d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3, %4}], [%5];" ::"r"(
                     smem_u32(dst_smem)),
                 "l"(map), "r"(x), "r"(y), "r"(z), "r"(smem_u32(bar))
                 : "memory");
}

__device__ __forceinline__ void tma_load_3d_hint(void* dst_smem, const CUtensorMap* map, int x, int y, int z, uint64_t* bar, uint64_t policy) {
    asm volatile("cp.async.bulk.tensor.3d.shared::cluster.global.mbarrier::complete_tx::bytes.L2::cache_hint [%0], [%1, {%2, %3, %4}], [%5], %6;" ::"r"(
                     smem_u32(dst_smem)),
                 "l"(map), "r"(x), "r"(y), "r"(z), "r"(smem_u32(bar)), "l"(policy)
                 : "memory");
}
__device__ __forceinline__ uint64_t l2_policy_evict_last() {
    uint64_t p;
    asm volatile("createpolicy.fractional.L2::evict_last.b64 %0, 1.0;" : "=l"(p));
    return p;
}
// Field stores of the step are streaming (st.global.cs: written once, next read a whole lattice later) and the staged planes
// are fetched with an L2 evict_last policy, so that the halo rows / columns a neighbouring CTA re-reads a moment later stay
// in L2 while the write stream passes through.  Measured at 512³ on one box, alternating (tools/exp_hints.sh,
// profiles/r01_exp_cache_hints.txt): plain/plain 1.0704 ms, streaming stores 1.0721, evict_last loads 1.0604, both 1.0575;
// evict_first loads 1.30 ms (the halo re-reads go to HBM).
__device__ __forceinline__ void stg4(float* p, const float4& v) { __stcs(reinterpret_cast<float4*>(p), v); }

struct FusedView {
    const float* in[6];   // Ex Ey Ez Hx Hy Hz of the current step (plane 0)
    float* out[6];        // the other buffer set
    const void* mat;
    const float4* table;
    int n_table, table_in_smem;
    int nx, ny, nz;       // nz = planes owned by this handle (z-slab); nx = row pitch
    int nxl;              // lattice cells of a row (≤ nx; the rest are pad cells, kernels.cuh) — read by the XPAD instantiations
    int z_lo, z_hi;       // planes updated by this launch
    int lo_open, hi_open; // a neighbouring slab exists below / above: ghost planes -1 / nz hold its old fields
    int late_issue;       // FDTD_B200_LATE_ISSUE=1: request the first planes after the forcing / probe scans (the A side of the A/B)
    float dx[3], inv_dx[3];
    // PUSH variant (boundary-plane launches of the peer-memory transport): the CTAs that own the slab's first / last plane
    // also store their results straight into the neighbouring GPUs' ghost planes over NVLink (peer-mapped pointers to the
    // start of that ghost plane in the neighbour's pending set; null = no neighbour on that side).
    float* push_down[5];  // E'x E'y E'z H'x H'y of plane 0      → the lower neighbour's ghost plane above its last plane
    float* push_up[2];    // E'x E'y of plane nz-1               → the upper neighbour's ghost plane below its plane 0
};

// CPML inside the fused step (template flag PML; SURVEY §8 f-2, formulas of kernels_pml.cuh).  The ψ recursion of an axis
// uses the very derivatives the curl needs, so the slabs cost their ψ traffic only.  ψ_H is read by the recomputed halo
// cells as well, so like the fields it is ping-pong buffered (gh_in → gh_out, written by the owner); ψ_E is touched by the
// owner of E' alone and advanced in place.  y and z slabs are x-contiguous (one float4 per quad and array); the few x-slab
// cells of a tile row are advanced one per lane (lane j: the row's j-th x-slab cell; its ψ is loaded before the phase's
// arithmetic and consumed after it) and handed to the cells' owners through shared memory.
struct FusedPmlView {
    int n[3];                  // PML coordinates per axis (0: none)
    const int* map[3];         // coordinate → index into the axis' slab list, -1 outside the slabs
    const float2* bc_h[3];     // (b, c) per slab index, H half-step
    const float2* bc_e[3];     // … E half-step
    const float* gh_in[3][2];  // ψ_H of the current set: G_ab, G_ac   layouts: x [z][y][k], y [z][k][x], z [k][y][x]
    float* gh_out[3][2];       // ψ_H of the pending set
    // z-slabs with a neighbour above: ψ_H of that neighbour's first plane (the recomputed plane nz), received into a
    // separate buffer; the pointers are biased so that the index of plane nz (axis z: of slab entry n[2]) lands in it
    const float* gh_ghost[3][2];
    float* ge[3][2];           // ψ_E
    // Dispatch order: the tile rows / plane chunks with slab work cost up to twice the others; launched last (high y, high z —
    // where a slab sits) they are the tail of the grid.  heavy[a] = {p, s}: block indices [0, p) and [s, count) along y (a = 1)
    // and z (a = 2) hold slab work and are mapped to the front: i → i (i < p), s + (i − p) (next count − s), else p + (i − p − (count − s)).
    int heavy[3][2];
};

__device__ __forceinline__ int heavy_first(int i, int p, int s, int count) {
    const int tail = count - s;
    return i < p ? i : (i < p + tail ? s + (i - p) : i - tail);
}

template <int BY>
struct FusedPmlSmem {
    static constexpr int XS = 32;     // x-slab cells of one tile row (both faces of a one-tile lattice: 2 × 16)
    static constexpr int ZMAX = 132;  // planes of one chunk + the recomputed plane above it
    float xg[2][2][BY][XS];           // ψ of the x slabs just advanced by lane j, for the cells' owners: [half-step][G_ab | G_ac][row][entry]
    float2 xbc[2][XS];                // (b, c) of each entry: [half-step]
    int xk[XS];                       // slab index of each entry
    int xl[XS];                       // tile column of each entry
    signed char xj[128];              // tile column → entry, -1 outside the slabs
    int wcnt[4];
    int yk[BY];
    float2 ybc[2][BY];
    int zk[ZMAX];
    float2 zbc[2][ZMAX];
    int zwin[2];                      // first and last chunk-relative plane (0 … n) that lies in a z slab; first > last: none
};

__device__ __forceinline__ void prefetch_l2(const void* p) { asm volatile("prefetch.global.L2 [%0];" ::"l"(p)); }

#ifdef FDTD_B200_CTA_TRACE
// Diagnostic build only (FDTD_B200_NVCC_EXTRA=-DFDTD_B200_CTA_TRACE): every CTA of a fused launch leaves (SM, start, end, tile) here
__device__ unsigned long long* g_cta_trace = nullptr;
__device__ __forceinline__ unsigned long long globaltimer_ns() { unsigned long long t; asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t)); return t; }
__device__ __forceinline__ unsigned smid() { unsigned s; asm volatile("mov.u32 %0, %%smid;" : "=r"(s)); return s; }
#endif

template <typename IdxT, int BX, int BY>
struct FusedLayout {
    static constexpr int NT = BX * BY;
    static constexpr int PE = 4 * BX + 8;                       // field row pitch: [left halo quad][4*BX cells][right halo quad]
    static constexpr int ME = 16 / static_cast<int>(sizeof(IdxT));  // extra index entries fetched to the right (16 bytes)
    static constexpr int PM = 4 * BX + ME;                      // index row pitch (entries)
    static constexpr int CORE_ROWS = BY - 1;                    // rows whose E' this CTA owns
    static constexpr int E_ROWS = BY + 1, H_ROWS = BY;          // old E: rows y0-1 .. y0+BY-1 ; H: rows y0 .. y0+BY-1
    // component strides are padded to 128 bytes: TMA tensor copies need 128-byte aligned shared-memory destinations
    static constexpr int E_COMP = (E_ROWS * PE + 31) / 32 * 32, H_COMP = (H_ROWS * PE + 31) / 32 * 32;
    static constexpr int E_BOX_BYTES = E_ROWS * PE * 4, H_BOX_BYTES = H_ROWS * PE * 4;
    static constexpr int M_BOX_BYTES = H_ROWS * PM * static_cast<int>(sizeof(IdxT));
    static constexpr int E_SLOT_BYTES = 3 * E_COMP * 4;
    static constexpr int MAT_BYTES = (M_BOX_BYTES + 127) / 128 * 128;
    static constexpr int H_SLOT_BYTES = 3 * H_COMP * 4 + MAT_BYTES;
    static constexpr int HB_BYTES = 3 * H_COMP * 4;
    static constexpr int NE = 3, NH = 2;                        // ring depths
    static constexpr size_t smem_bytes(int n_table_in_smem, bool pml = false) {
        return 128 + static_cast<size_t>(NE) * E_SLOT_BYTES + static_cast<size_t>(NH) * H_SLOT_BYTES + 2 * static_cast<size_t>(HB_BYTES) +
               static_cast<size_t>(n_table_in_smem) * sizeof(float4) + (pml ? sizeof(FusedPmlSmem<BY>) : 0);
    }
};

// XPAD: rows carry pad cells (nxl < nx, lattices whose nx is not a multiple of 16): the x+1 Neumann test of E' is made per cell
// against nxl instead of per quad against the pitch; everything else is untouched (pad cells have all-zero coefficients).
template <int EXACT, typename IdxT, int BX, int BY, bool PUSH = false, bool PML = false, bool XPAD = false>
__global__ void __launch_bounds__(BX * BY, 2)
k_step_fused(const __grid_constant__ FusedMaps M, const FusedView V, const ForcingView F, const ProbeView P, const int zc, const FusedPmlView Q) {
    using LO = FusedLayout<IdxT, BX, BY>;
    constexpr int NT = LO::NT, PE = LO::PE, PM = LO::PM;
    constexpr int E_COMP = LO::E_COMP, H_COMP = LO::H_COMP;
    static_assert(LO::E_SLOT_BYTES % 128 == 0 && LO::H_SLOT_BYTES % 128 == 0 && LO::HB_BYTES % 128 == 0, "128-byte aligned slots");
    // BX = 32, BY = 8: one warp per thread row, tile 128 cells × 7 core rows.  BX = 16, BY = 16 (plain instantiation only): two thread
    // rows per warp, tile 64 cells × 15 core rows — for lattices whose last 128-cell tile column would be at most half full.  The last
    // warp issues the copies and owns the right-column cells either way; with two rows per warp it also owns E' of row BY − 2.
    static_assert((BX == 32 || BX == 16) && (BX * BY) % 32 == 0 && BY <= 32, "whole warps; the producer warp's lanes cover the BY right-column cells");
    static_assert(BX == 32 || !PML, "the CPML blocks map lane j to the row's j-th x-slab cell: one thread row per warp");
    constexpr int WARPS = BX * BY / 32;

    griddep_launch_dependents();
    extern __shared__ __align__(128) unsigned char smem_raw[];
    uint64_t* fullE = reinterpret_cast<uint64_t*>(smem_raw);       // 3 barriers: TMA data of an E ring slot has landed
    uint64_t* fullH = fullE + LO::NE;                              // 2 barriers: … of an H ring slot
    uint64_t* emptyE = fullH + LO::NH;                             // 3 barriers: all BY warps have read the slot into registers
    uint64_t* emptyH = emptyE + LO::NE;                            // 2 barriers
    // bytes 80 … 127 of the header: which planes of this CTA's march hold a source / a probe (bit k: plane za + k; a march is
    // at most 128 planes + the recomputed one above: 5 words each)
    unsigned* src_planes = reinterpret_cast<unsigned*>(smem_raw + 80);
    unsigned* probe_planes = src_planes + 5;
    unsigned char* e_slots = smem_raw + 128;
    unsigned char* h_slots = e_slots + LO::NE * LO::E_SLOT_BYTES;
    float* hb = reinterpret_cast<float*>(h_slots + LO::NH * LO::H_SLOT_BYTES);   // 2 × 3 × H_COMP
    float4* s_tab = reinterpret_cast<float4*>(reinterpret_cast<unsigned char*>(hb) + 2 * LO::HB_BYTES);

    // The PML instantiation is bound by instruction issue (≈ 1.5 × the warp instructions of the plain one), so it sheds what it
    // can: no run-time test of where the coefficient table lives for 8-bit indices (≤ 256 rows are always in shared memory) and
    // a quad loop without the forcing lookup where the CTA holds no source.  The plain instantiation waits on HBM and measured
    // 2 % SLOWER with the same trimming (123.9 vs 126.6 Gcells/s at 512³, same box), so it keeps the code it was tuned with.
    constexpr bool LEAN = PML;
    auto coef = [&](int m) -> float4 {
        if (LEAN && sizeof(IdxT) == 1) return s_tab[m];
        return V.table_in_smem ? s_tab[m] : __ldg(V.table + m);
    };
    const int tid = threadIdx.y * BX + threadIdx.x;
    const unsigned lane = tid & 31u;
    const int x0 = blockIdx.x * (4 * BX);
    const int y0 = (PML ? heavy_first(blockIdx.y, Q.heavy[1][0], Q.heavy[1][1], gridDim.y) : static_cast<int>(blockIdx.y)) * LO::CORE_ROWS;
    const int za = V.z_lo + (PML ? heavy_first(blockIdx.z, Q.heavy[2][0], Q.heavy[2][1], gridDim.z) : static_cast<int>(blockIdx.z)) * zc;
    const int zb = min(za + zc, V.z_hi);
    const int n = zb - za;
    const size_t sy = static_cast<size_t>(V.nx);
    const size_t sz = sy * static_cast<size_t>(V.ny);
    const bool right = x0 + 4 * BX < V.nx;
    const bool below = (za > 0) || (V.lo_open != 0);                // old-E plane za-1 exists (own plane or ghost)
    const int last_h_plane = (zb < V.nz || V.hi_open != 0) ? zb : zb - 1;   // H'/old-E planes za..last_h_plane are needed
    const int n_h = last_h_plane - za + 1;                          // number of H-plane iterations with a phase A

#ifdef FDTD_B200_CTA_TRACE
    const size_t trace_at = 4 * (blockIdx.x + static_cast<size_t>(gridDim.x) * (blockIdx.y + static_cast<size_t>(gridDim.y) * blockIdx.z));
    if (tid == 0 && g_cta_trace) {
        g_cta_trace[trace_at] = smid(); g_cta_trace[trace_at + 1] = globaltimer_ns();
        g_cta_trace[trace_at + 3] = (static_cast<unsigned long long>(x0) << 40) | (static_cast<unsigned long long>(y0) << 20) | static_cast<unsigned long long>(za);
    }
#endif
    const uint64_t ld_policy = l2_policy_evict_last();
    // ---- producer: one lane of the last warp issues 7 tensor copies per plane ------------------------------------
    auto issue_e = [&](int plane, int slot) {
        if (lane == 0) {
            uint64_t* bar = fullE + slot;
            mbar_arrive_expect_tx(bar, 3u * LO::E_BOX_BYTES);
            float* s = reinterpret_cast<float*>(e_slots + static_cast<size_t>(slot) * LO::E_SLOT_BYTES);
#pragma unroll
            for (int c = 0; c < 3; ++c) tma_load_3d_hint(s + c * E_COMP, &M.e[c], x0 - 4, y0 - 1, plane + 1, bar, ld_policy);
        }
    };
    auto issue_h = [&](int plane, int slot) {
        if (lane == 0) {
            uint64_t* bar = fullH + slot;
            mbar_arrive_expect_tx(bar, 3u * LO::H_BOX_BYTES + LO::M_BOX_BYTES);
            unsigned char* sl = h_slots + static_cast<size_t>(slot) * LO::H_SLOT_BYTES;
            float* s = reinterpret_cast<float*>(sl);
#pragma unroll
            for (int c = 0; c < 3; ++c) tma_load_3d_hint(s + c * H_COMP, &M.h[c], x0 - 4, y0, plane + 1, bar, ld_policy);
            tma_load_3d(sl + 3 * H_COMP * 4, &M.mat, x0, y0, plane, bar);
        }
    };
    // E ring index of plane z: e = z - (za - 1); H ring index: k = z - za
    const bool producer = tid >= NT - 32;                           // the last warp: it has no E' work (top thread row)
    auto first_issue = [&]() {
        if (producer && n > 0) {
        if (below) issue_e(za - 1, 0);
        else if (lane == 0) mbar_arrive_expect_tx(fullE + 0, 0u);   // keep slot 0's phase sequence uniform (phase 0 completes empty)
        issue_e(za, 1);
        if (n_h > 1) issue_e(za + 1, 2);
        issue_h(za, 0);
        if (n_h > 1) issue_h(za + 1, 1);
        }
    };
    // Lane 0 of the producer warp initialises the barriers and — in the plain instantiation — requests the first planes at once,
    // before the coefficient table is staged and before the per-CTA forcing / probe scans, so that those overlap the copies'
    // latency (the other warps see the barriers after the __syncthreads below).  Measured against issuing after the scans,
    // same box, alternating (profiles/r01_exp_early_issue.txt): 512³ 129.8 vs 129.1 Gcells/s, 256³ 112.3 vs 109.3, 384³ 125.2 vs
    // 122.5, 1024×1024×128 127.7 vs 126.8, 4096×4096×1 68.0 vs 52.0; issuing after the first barrier instead is equal within
    // 0.3 %.  The PML instantiation measured 0.8 % slower with the early issue and keeps the late one.
    constexpr bool EARLY = !PML;
    if (tid < 10) src_planes[tid] = 0u;   // (both masks; ordered before the scans by the __syncthreads below)
    if (tid == NT - 32) {
        for (int s = 0; s < LO::NE; ++s) mbar_init(fullE + s, 1);
        for (int s = 0; s < LO::NH; ++s) mbar_init(fullH + s, 1);
        for (int s = 0; s < LO::NE; ++s) mbar_init(emptyE + s, WARPS);
        for (int s = 0; s < LO::NH; ++s) mbar_init(emptyH + s, WARPS);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    }
    griddep_wait();   // the previous step's grid is complete: its fields may be read, its inputs overwritten
    if (EARLY && !V.late_issue) first_issue();
    if (V.table_in_smem) {
        for (int t = tid; t < V.n_table; t += NT) s_tab[t] = V.table[t];
    }
    // ---- CPML: which of the tile's columns, rows and planes lie in a slab, and their (b, c) ------------------------
    using PS = FusedPmlSmem<BY>;
    PS& S = *reinterpret_cast<PS*>(s_tab + (V.table_in_smem ? V.n_table : 0));
    int x_rank = 0, x_k = -1;
    if (PML) {
        static_assert(!PML || (BY >= 4 && 4 * BX == 128), "the first four warps scan the tile's 128 columns");
        if (tid < 4 * BX) {
            const int gx = x0 + tid;
            if (Q.map[0] != nullptr && gx < V.nx) x_k = __ldg(Q.map[0] + gx);
            const unsigned m = __ballot_sync(0xffffffffu, x_k >= 0);
            x_rank = __popc(m & ((1u << lane) - 1u));
            if (lane == 0) S.wcnt[tid >> 5] = __popc(m);
        }
        if (tid < BY) {
            const int gy = y0 + tid;
            const int ky = (Q.map[1] != nullptr && gy < V.ny) ? __ldg(Q.map[1] + gy) : -1;
            S.yk[tid] = ky;
            if (ky >= 0) { S.ybc[0][tid] = __ldg(Q.bc_h[1] + ky); S.ybc[1][tid] = __ldg(Q.bc_e[1] + ky); }
        }
        for (int t = tid; t <= n && t < PS::ZMAX; t += NT) {
            const int z = za + t;
            const int kz = (Q.map[2] != nullptr && (z < V.nz || (z == V.nz && V.hi_open))) ? __ldg(Q.map[2] + z) : -1;   // map[nz]: the neighbour's plane
            S.zk[t] = kz;
            if (kz >= 0) { S.zbc[0][t] = __ldg(Q.bc_h[2] + kz); S.zbc[1][t] = __ldg(Q.bc_e[2] + kz); }
        }
        if (tid == 0) { S.zwin[0] = PS::ZMAX; S.zwin[1] = -1; }
    }
    __syncthreads();
    if (PML) {
        for (int t = tid; t <= n && t < PS::ZMAX; t += NT)
            if (S.zk[t] >= 0) { atomicMin(&S.zwin[0], t); atomicMax(&S.zwin[1], t); }
    }
    // the recomputed column right of the tile (x0 + 128): its x-slab entry and (b, c), if a slab crosses the tile boundary there
    int xr_k = -1;
    float2 xr_bc = make_float2(0.f, 0.f);
    if (PML && producer && Q.map[0] != nullptr && x0 + 4 * BX < V.nx) {
        xr_k = __ldg(Q.map[0] + x0 + 4 * BX);
        if (xr_k >= 0) xr_bc = __ldg(Q.bc_h[0] + xr_k);
    }
    int nxs = 0;          // x-slab cells of a tile row (entries of the hand-over rows)
    int xjp = -1;         // entries of the own quad's four cells, one signed byte each
    // Most CTAs of a lattice touch no y or z slab at all: everything those slabs need per iteration (tests, prefetches, the
    // blocks themselves) sits behind these CTA-uniform flags, so an interior CTA pays one branch per phase for them.
    bool cta_y = false;   // one of the tile's rows lies in a y slab
    int zw_lo = 1 << 30, zw_hi = -1;   // chunk-relative planes (0 … n) that lie in a z slab: inside [zw_lo, zw_hi]
    if (PML) {
        nxs = S.wcnt[0] + S.wcnt[1] + S.wcnt[2] + S.wcnt[3];
        if (tid < 4 * BX) {
            int off = 0;
            for (int w = 0; w < (tid >> 5); ++w) off += S.wcnt[w];
            const int j = x_k >= 0 ? off + x_rank : -1;
            S.xj[tid] = static_cast<signed char>(j);
            if (j >= 0 && j < PS::XS) { S.xk[j] = x_k; S.xl[j] = tid; S.xbc[0][j] = __ldg(Q.bc_h[0] + x_k); S.xbc[1][j] = __ldg(Q.bc_e[0] + x_k); }
        }
        __syncthreads();
        xjp = *reinterpret_cast<const int*>(&S.xj[4 * threadIdx.x]);
#pragma unroll
        for (int r = 0; r < BY; ++r) cta_y = cta_y || S.yk[r] >= 0;
        zw_lo = S.zwin[0]; zw_hi = S.zwin[1];
    }

    const long long sz_ll = static_cast<long long>(V.nx) * V.ny;
    // forcing may hit the recomputed halo as well: test the extended box
    bool has_src = false, has_probe = false;
    // … and which planes: a CTA that holds a source or a probe pays for the per-cell lookups on that plane only.  (Measured with
    // per-CTA time stamps, 256³ in one wave of 64-plane marches: the source's CTA ran 165 µs and the probes' CTAs 137–144 µs against a
    // median of 128 µs when every plane of their march did the lookups — and the slowest CTA of the wave is the kernel's time.)
    if (F.n) has_src = __syncthreads_or(mark_planes_in_box(F.cells, F.n, tid, NT, x0, x0 + 4 * BX + 4, y0, y0 + BY, za, zb + 1, V.nx, sz_ll, src_planes)) != 0;
    if (P.n) has_probe = __syncthreads_or(mark_planes_in_box(P.cells, P.n, tid, NT, x0, x0 + 4 * BX, y0, y0 + LO::CORE_ROWS, za, zb, V.nx, sz_ll, probe_planes)) != 0;
    if (!EARLY || V.late_issue) first_issue();
    if (n <= 0) return;

    const int tx = threadIdx.x, ty = threadIdx.y;
    const int x = x0 + 4 * tx, y = y0 + ty;
    const bool act = (x < V.nx) && (y < V.ny);
    const bool core = act && (ty < LO::CORE_ROWS);                  // this thread owns E' of its quad
    // phase B is entered warp by warp: a warp whose first thread row is a core row.  With one row per warp that is the thread's own
    // `ty < CORE_ROWS`; with two, the last warp's upper half (row BY − 1) walks through phase B as a shadow of row BY − 2 — tyb keeps its
    // shared-memory reads inside the tiles — and stores nothing (`core`).
    const bool warp_core = ((tid & ~31) / BX) < LO::CORE_ROWS;
    const int tyb = (BX == 32) ? ty : min(ty, LO::CORE_ROWS - 1);
    const int nx_cells = XPAD ? V.nxl : V.nx;                       // lattice cells of a row
    const int xlim = nx_cells - 1 - x;                              // XPAD: cell q of the quad has an x+1 neighbour iff q < xlim

    float4 hxo = make_float4(0.f, 0.f, 0.f, 0.f), hyo = hxo, hzo = hxo;   // H'(z_k) of the own quad
    int mk_prev[4] = {0, 0, 0, 0};

    // CPML, x slabs: lane j owns the row's j-th x-slab cell for the whole march — its slab index and tile column never change
    // (32-bit indices: the host keeps the recursion in the kernel only while an x-slab array has fewer than 2^31 entries)
    const bool xlane = PML && static_cast<int>(lane) < nxs && y < V.ny;
    int xl_j = 0;
    unsigned xq_row = 0;  // index of (plane 0, row y, entry j) in the x-slab arrays [z][y][k]; a plane is ny·n[0] entries
    if (xlane) { xl_j = S.xl[lane]; xq_row = static_cast<unsigned>(y) * static_cast<unsigned>(Q.n[0]) + static_cast<unsigned>(S.xk[lane]); }
    const unsigned xq_plane = PML ? static_cast<unsigned>(V.ny) * static_cast<unsigned>(Q.n[0]) : 0u;

    // does chunk-relative plane i hold a source / a probe?  Read behind the loop-invariant has_src / has_probe only, so that the
    // march of a CTA without any (all but a handful) stays the loop the compiler unswitches
    auto plane_bit = [](const unsigned* mask, int i) -> bool { return (mask[i >> 5] >> (i & 31)) & 1u; };
    for (int k = 0; k <= n; ++k) {
        const int zk = za + k;
        const bool has_a = k < n_h;
        const float* e_cur = nullptr;
        float4 hxn = hxo, hyn = hyo, hzn = hzo;
        int mk_cur[4] = {0, 0, 0, 0};
        const bool zwin_a = PML && k >= zw_lo && k <= zw_hi;           // plane z_k (phase A) lies in a z slab — CTA-uniform
        const bool zwin_b = PML && k - 1 >= zw_lo && k - 1 <= zw_hi;   // plane z_k - 1 (phase B) does
        if (PML && (nxs || cta_y || (k + 1 >= zw_lo && k <= zw_hi)) && act) {
            // the ψ of the next iteration's cells (plain loads there): pull the lines into L2 now
            const int kk = k + 1;
            const bool nh = kk < n_h && za + kk < V.nz, ne = ty < LO::CORE_ROWS && kk <= n;
            if (nxs) {
                if (xlane) {
                    const unsigned q0 = static_cast<unsigned>(za + kk) * xq_plane + xq_row;
                    if (nh) { prefetch_l2(Q.gh_in[0][0] + q0); prefetch_l2(Q.gh_in[0][1] + q0); }
                    if (ne) { prefetch_l2(Q.ge[0][0] + (q0 - xq_plane)); prefetch_l2(Q.ge[0][1] + (q0 - xq_plane)); }
                }
            }
            if (cta_y) {
                const int pky = S.yk[ty];
                if (pky >= 0) {
                    const size_t q0 = (static_cast<size_t>(za + kk - 1) * Q.n[1] + pky) * sy + x, dq = static_cast<size_t>(Q.n[1]) * sy;
                    if (nh) { prefetch_l2(Q.gh_in[1][0] + q0 + dq); prefetch_l2(Q.gh_in[1][1] + q0 + dq); }
                    if (ne) { prefetch_l2(Q.ge[1][0] + q0); prefetch_l2(Q.ge[1][1] + q0); }
                }
            }
            if (k + 1 >= zw_lo && k <= zw_hi) {
                const int kzh = nh ? S.zk[kk] : -1, kze = ne ? S.zk[kk - 1] : -1;
                if (kzh >= 0) {
                    const size_t q0 = (static_cast<size_t>(kzh) * V.ny + y) * sy + x;
                    prefetch_l2(Q.gh_in[2][0] + q0); prefetch_l2(Q.gh_in[2][1] + q0);
                }
                if (kze >= 0) {
                    const size_t q0 = (static_cast<size_t>(kze) * V.ny + y) * sy + x;
                    prefetch_l2(Q.ge[2][0] + q0); prefetch_l2(Q.ge[2][1] + q0);
                }
            }
        }
        if (has_a) {
            const int e_idx = k + 1;
            // CPML, x slabs: lane j's ψ_H of plane z_k — requested before the wait for the planes, used after the quad loop
            float xp0 = 0.f, xp1 = 0.f;
            if (PML && nxs) {
                if (xlane) {
                    const unsigned qi = static_cast<unsigned>(zk) * xq_plane + xq_row;
                    const bool ghost = zk == V.nz;   // the plane above a z-slab: the upper neighbour's ψ_H, read-only here
                    xp0 = __ldg((ghost ? Q.gh_ghost[0][0] : Q.gh_in[0][0]) + qi);
                    xp1 = __ldg((ghost ? Q.gh_ghost[0][1] : Q.gh_in[0][1]) + qi);
                }
            }
            mbar_wait(fullE + (e_idx % LO::NE), static_cast<uint32_t>((e_idx / LO::NE) & 1));
            mbar_wait(fullH + (k % LO::NH), static_cast<uint32_t>((k / LO::NH) & 1));
            if (k == 0 && below) mbar_wait(fullE + 0, 0u);
            e_cur = reinterpret_cast<const float*>(e_slots + static_cast<size_t>(e_idx % LO::NE) * LO::E_SLOT_BYTES);
            const float* e_low = reinterpret_cast<const float*>(e_slots + static_cast<size_t>(k % LO::NE) * LO::E_SLOT_BYTES);
            const unsigned char* hsl = h_slots + static_cast<size_t>(k % LO::NH) * LO::H_SLOT_BYTES;
            const float* h_old = reinterpret_cast<const float*>(hsl);
            const IdxT* s_mat = reinterpret_cast<const IdxT*>(hsl + 3 * H_COMP * 4);
            float* hb_w = hb + static_cast<size_t>(k & 1) * 3 * H_COMP;
            const bool z_in = (zk > 0) || (V.lo_open != 0);

            // ---- phase A: H'(z_k) on the CTA's BY rows -----------------------------------------------------------
            {
                const int xq = tx, r = ty;
                const bool ok = act;
                const int c0 = 4 + 4 * xq;
                const int gx = x0 + 4 * xq, gy = y0 + r;
                const float* pe = e_cur + (r + 1) * PE + c0;
                const float4 ex = ld4(pe), ey = ld4(pe + E_COMP), ez = ld4(pe + 2 * E_COMP);
                const float4 ex_ym = ld4(pe - PE), ez_ym = ld4(pe + 2 * E_COMP - PE);
                const float ey_l = pe[E_COMP - 1], ez_l = pe[2 * E_COMP - 1];
                const float* pl = e_low + (r + 1) * PE + c0;
                float4 ex_zm = ld4(pl), ey_zm = ld4(pl + E_COMP);
                if (!z_in) ex_zm = ey_zm = make_float4(0.f, 0.f, 0.f, 0.f);   // no plane below the lattice: that ring slot was never filled, and the masked reciprocal needs finite operands
                const float* ph = h_old + r * PE + c0;
                float4 hx = ld4(ph), hy = ld4(ph + H_COMP), hz = ld4(ph + 2 * H_COMP);
                int mk[4];
                Idx4<IdxT>::load(s_mat, static_cast<size_t>(r) * PM + 4 * xq, mk);
                if (!ok) { mk[0] = mk[1] = mk[2] = mk[3] = 0; }
                const bool y_in = gy > 0, x_in = gx > 0;
                const float idy = y_in ? V.inv_dx[1] : 0.f, idz = z_in ? V.inv_dx[2] : 0.f;
                long long cell = 0;
                int s0 = 0, cnt = 0;
                if (!LEAN) cell = static_cast<long long>(zk) * sz_ll + static_cast<long long>(gy) * V.nx + gx;
                if (has_src && plane_bit(src_planes, k)) {
                    if (LEAN) cell = static_cast<long long>(zk) * sz_ll + static_cast<long long>(gy) * V.nx + gx;
                    s0 = lower_bound_ll(F.cells, F.n, cell);
                    while (s0 + cnt < F.n && __ldg(F.cells + s0 + cnt) < cell + 4) ++cnt;
                }
                const bool own_h = core && k < n;
                const float exv[4] = {ex.x, ex.y, ex.z, ex.w};
                const float eyv[4] = {ey.x, ey.y, ey.z, ey.w};
                const float ezv[4] = {ez.x, ez.y, ez.z, ez.w};
                const float exym[4] = {ex_ym.x, ex_ym.y, ex_ym.z, ex_ym.w};
                const float ezym[4] = {ez_ym.x, ez_ym.y, ez_ym.z, ez_ym.w};
                const float exzm[4] = {ex_zm.x, ex_zm.y, ex_zm.z, ex_zm.w};
                const float eyzm[4] = {ey_zm.x, ey_zm.y, ey_zm.z, ey_zm.w};
                const float eyxm[4] = {ey_l, ey.x, ey.y, ey.z};
                const float ezxm[4] = {ez_l, ez.x, ez.y, ez.z};
                float hxv[4] = {hx.x, hx.y, hx.z, hx.w};
                float hyv[4] = {hy.x, hy.y, hy.z, hy.w};
                float hzv[4] = {hz.x, hz.y, hz.z, hz.w};
                // the quad's four cells; instantiated with and without the forcing lookup (one CTA in thousands holds a source)
                auto h_quad = [&](auto with_src) {
#pragma unroll
                for (int q = 0; q < 4; ++q) {
                    const bool xq_in = (q > 0) || x_in;
                    const float dez_dy = deriv_masked<EXACT>(ezv[q], ezym[q], V.dx[1], idy, y_in);
                    const float dex_dy = deriv_masked<EXACT>(exv[q], exym[q], V.dx[1], idy, y_in);
                    const float dey_dz = deriv_masked<EXACT>(eyv[q], eyzm[q], V.dx[2], idz, z_in);
                    const float dex_dz = deriv_masked<EXACT>(exv[q], exzm[q], V.dx[2], idz, z_in);
                    const float dez_dx = xq_in ? deriv<EXACT>(ezv[q], ezxm[q], V.dx[0], V.inv_dx[0]) : 0.f;
                    const float dey_dx = xq_in ? deriv<EXACT>(eyv[q], eyxm[q], V.dx[0], V.inv_dx[0]) : 0.f;
                    const float cx = dez_dy - dey_dz;
                    const float cy = dex_dz - dez_dx;
                    const float cz = dey_dx - dex_dy;
                    float mx = 0.f, my = 0.f, mz = 0.f;
                    if (decltype(with_src)::value && cnt) { const float3 v = forcing_lookup(F, s0, cnt, cell + q, 3); mx = v.x; my = v.y; mz = v.z; }
                    const float4 c = coef(mk[q]);
                    hxv[q] = leap<EXACT>(c.z, hxv[q], c.w, -cx, mx);
                    hyv[q] = leap<EXACT>(c.z, hyv[q], c.w, -cy, my);
                    hzv[q] = leap<EXACT>(c.z, hzv[q], c.w, -cz, mz);
                }
                };
                if (!LEAN || has_src) h_quad(std::true_type{}); else h_quad(std::false_type{});   // (loop-invariant choice; on planes without a source cnt is 0)
                if (PML && nxs) {
                    // x slabs (k_pml_axis<0, 0>): lane j advances ψ_H of the row's j-th x-slab cell from ∂Ey/∂x, ∂Ez/∂x of the staged
                    // E plane, stores it if this CTA owns the row, and hands it to the cell's owner, who adds H_y += d_b·G_xz,
                    // H_z −= d_b·G_xy — outside the unrolled quad loop, so that the 28 lanes of a row without x-slab cells pay
                    // one block of predicated instructions instead of four
                    if (xlane) {
                        const float* pj = e_cur + (r + 1) * PE + 4 + xl_j;
                        const bool in = x0 + xl_j > 0;
                        const float dey = in ? deriv<EXACT>(pj[E_COMP], pj[E_COMP - 1], V.dx[0], V.inv_dx[0]) : 0.f;
                        const float dez = in ? deriv<EXACT>(pj[2 * E_COMP], pj[2 * E_COMP - 1], V.dx[0], V.inv_dx[0]) : 0.f;
                        const float2 bc = S.xbc[0][lane];
                        const float g0 = fmaf(bc.x, xp0, bc.y * dey);
                        const float g1 = fmaf(bc.x, xp1, bc.y * dez);
                        S.xg[0][0][r][lane] = g0; S.xg[0][1][r][lane] = g1;
                        if (k < n && r < LO::CORE_ROWS) {
                            const unsigned qi = static_cast<unsigned>(zk) * xq_plane + xq_row;
                            Q.gh_out[0][0][qi] = g0; Q.gh_out[0][1][qi] = g1;
                        }
                    }
                    __syncwarp();
                    if (xjp != -1 && ok) {
#pragma unroll
                        for (int q = 0; q < 4; ++q) {
                            const int j = (xjp << (24 - 8 * q)) >> 24;
                            if (j >= 0) {
                                const float w = coef(mk[q]).w;
                                hyv[q] = fmaf(w, S.xg[0][1][r][j], hyv[q]);
                                hzv[q] = fmaf(-w, S.xg[0][0][r][j], hzv[q]);
                            }
                        }
                    }
                }
                if (PML && (cta_y || zwin_a)) {
                    // y and z slabs (whole rows / planes: warp- and CTA-uniform): G ← b·G + c·∂E/∂x_a from the staged planes
                    // again, H_b += d_b·G_ac, H_c −= d_b·G_ab; y before z as the passes run (k_pml_axis<0, 1>, <0, 2>)
                    const int pky = ok ? S.yk[r] : -1, pkz = ok ? S.zk[k] : -1;
                    const bool ghost_h = zk == V.nz;   // the plane above a z-slab: the upper neighbour's ψ_H, read-only here
                    if (pky >= 0) {
                        const size_t qy = (static_cast<size_t>(zk) * Q.n[1] + pky) * sy + gx;
                        const float4 a = ld4((ghost_h ? Q.gh_ghost[1][0] : Q.gh_in[1][0]) + qy), b = ld4((ghost_h ? Q.gh_ghost[1][1] : Q.gh_in[1][1]) + qy);
                        const float2 bc = S.ybc[0][r];
                        float g0[4] = {a.x, a.y, a.z, a.w}, g1[4] = {b.x, b.y, b.z, b.w};
#pragma unroll
                        for (int q = 0; q < 4; ++q) {
                            const float w = coef(mk[q]).w;
                            g0[q] = fmaf(bc.x, g0[q], bc.y * deriv_masked<EXACT>(pe[2 * E_COMP + q], pe[2 * E_COMP - PE + q], V.dx[1], idy, y_in));
                            g1[q] = fmaf(bc.x, g1[q], bc.y * deriv_masked<EXACT>(pe[q], pe[q - PE], V.dx[1], idy, y_in));
                            hzv[q] = fmaf(w, g1[q], hzv[q]);
                            hxv[q] = fmaf(-w, g0[q], hxv[q]);
                        }
                        if (own_h) {
                            st4(Q.gh_out[1][0] + qy, make_float4(g0[0], g0[1], g0[2], g0[3]));
                            st4(Q.gh_out[1][1] + qy, make_float4(g1[0], g1[1], g1[2], g1[3]));
                        }
                    }
                    if (pkz >= 0) {
                        const size_t qz = (static_cast<size_t>(pkz) * V.ny + gy) * sy + gx;
                        const float4 a = ld4((ghost_h ? Q.gh_ghost[2][0] : Q.gh_in[2][0]) + qz), b = ld4((ghost_h ? Q.gh_ghost[2][1] : Q.gh_in[2][1]) + qz);
                        const float2 bc = S.zbc[0][k];
                        float g0[4] = {a.x, a.y, a.z, a.w}, g1[4] = {b.x, b.y, b.z, b.w};
#pragma unroll
                        for (int q = 0; q < 4; ++q) {
                            const float w = coef(mk[q]).w;
                            const float lo_x = z_in ? pl[q] : 0.f, lo_y = z_in ? pl[E_COMP + q] : 0.f;   // as ex_zm, ey_zm above
                            g0[q] = fmaf(bc.x, g0[q], bc.y * deriv_masked<EXACT>(pe[q], lo_x, V.dx[2], idz, z_in));
                            g1[q] = fmaf(bc.x, g1[q], bc.y * deriv_masked<EXACT>(pe[E_COMP + q], lo_y, V.dx[2], idz, z_in));
                            hxv[q] = fmaf(w, g1[q], hxv[q]);
                            hyv[q] = fmaf(-w, g0[q], hyv[q]);
                        }
                        if (own_h) {
                            st4(Q.gh_out[2][0] + qz, make_float4(g0[0], g0[1], g0[2], g0[3]));
                            st4(Q.gh_out[2][1] + qz, make_float4(g1[0], g1[1], g1[2], g1[3]));
                        }
                    }
                }
                hx = make_float4(hxv[0], hxv[1], hxv[2], hxv[3]);
                hy = make_float4(hyv[0], hyv[1], hyv[2], hyv[3]);
                hz = make_float4(hzv[0], hzv[1], hzv[2], hzv[3]);
                float* pw = hb_w + r * PE + c0;
                st4(pw, hx); st4(pw + H_COMP, hy); st4(pw + 2 * H_COMP, hz);
                {
                    hxn = hx; hyn = hy; hzn = hz;
                    mk_cur[0] = mk[0]; mk_cur[1] = mk[1]; mk_cur[2] = mk[2]; mk_cur[3] = mk[3];
                    if (core && k < n) {                          // this CTA owns H' of its core rows and planes
                        const size_t gi = static_cast<size_t>(zk) * sz + static_cast<size_t>(gy) * sy + static_cast<size_t>(gx);
                        stg4(V.out[3] + gi, hx); stg4(V.out[4] + gi, hy); stg4(V.out[5] + gi, hz);
                        if (PUSH && zk == 0 && V.push_down[3] != nullptr) {
                            const size_t pi = static_cast<size_t>(gy) * sy + static_cast<size_t>(gx);
                            st4(V.push_down[3] + pi, hx); st4(V.push_down[4] + pi, hy);
                        }
                    }
                }
            }
            // H'y, H'z of the single cell to the right of each row (x1 = x0 + 4*BX): phase B's d/dx of the row's last cell.
            // The top thread row's warp (no E' work in phase B) computes all BY of them, one row per lane.
            if (right && producer && lane < BY) {
                const int r = static_cast<int>(lane), cR = 4 + 4 * BX;
                const int gy = y0 + r, gx = x0 + 4 * BX;
                const float* pe = e_cur + (r + 1) * PE + cR;
                const float ex = pe[0], ex_ym = pe[-PE], ey = pe[E_COMP], ey_l = pe[E_COMP - 1], ez = pe[2 * E_COMP], ez_l = pe[2 * E_COMP - 1];
                const float ex_zm = z_in ? e_low[(r + 1) * PE + cR] : 0.f;
                const float hy = h_old[H_COMP + r * PE + cR], hz = h_old[2 * H_COMP + r * PE + cR];
                int m1 = static_cast<int>(s_mat[static_cast<size_t>(r) * PM + 4 * BX]);
                if (gy >= V.ny) m1 = 0;
                const bool y_in = gy > 0;
                const float dex_dz = z_in ? deriv<EXACT>(ex, ex_zm, V.dx[2], V.inv_dx[2]) : 0.f;
                const float dez_dx = deriv<EXACT>(ez, ez_l, V.dx[0], V.inv_dx[0]);
                const float dey_dx = deriv<EXACT>(ey, ey_l, V.dx[0], V.inv_dx[0]);
                const float dex_dy = y_in ? deriv<EXACT>(ex, ex_ym, V.dx[1], V.inv_dx[1]) : 0.f;
                const float cy = dex_dz - dez_dx;
                const float cz = dey_dx - dex_dy;
                float my = 0.f, mz = 0.f;
                if (has_src && plane_bit(src_planes, k)) {
                    const long long cell = static_cast<long long>(zk) * sz_ll + static_cast<long long>(gy) * V.nx + gx;
                    const int s1 = lower_bound_ll(F.cells, F.n, cell);
                    if (s1 < F.n && __ldg(F.cells + s1) == cell) { my = __ldg(F.vals + 6 * s1 + 4); mz = __ldg(F.vals + 6 * s1 + 5); }
                }
                const float4 c = coef(m1);
                float hy1 = leap<EXACT>(c.z, hy, c.w, -cy, my);
                float hz1 = leap<EXACT>(c.z, hz, c.w, -cz, mz);
                if (PML && gy < V.ny) {
                    // recomputed, the owner stores them: if the column right of the tile lies in an x slab (a slab that crosses
                    // a tile boundary) its ψ_H is advanced from ∂Ey/∂x, ∂Ez/∂x (k_pml_axis<0, 0>: H_y += d_b·G_xz, H_z −= d_b·G_xy);
                    // then H'z takes the y slab's G_yx (∂Ex/∂y), H'y the z slab's G_zx (∂Ex/∂z)
                    if (xr_k >= 0) {
                        const size_t qi = (static_cast<size_t>(zk) * V.ny + gy) * Q.n[0] + xr_k;
                        const bool ghost = zk == V.nz;
                        const float g0 = fmaf(xr_bc.x, (ghost ? Q.gh_ghost[0][0] : Q.gh_in[0][0])[qi], xr_bc.y * dey_dx);
                        const float g1 = fmaf(xr_bc.x, (ghost ? Q.gh_ghost[0][1] : Q.gh_in[0][1])[qi], xr_bc.y * dez_dx);
                        hy1 = fmaf(c.w, g1, hy1);
                        hz1 = fmaf(-c.w, g0, hz1);
                    }
                    const int ky = S.yk[r], kz = S.zk[k];
                    if (ky >= 0) {
                        const float2 bc = S.ybc[0][r];
                        const float g1 = fmaf(bc.x, (zk == V.nz ? Q.gh_ghost[1][1] : Q.gh_in[1][1])[(static_cast<size_t>(zk) * Q.n[1] + ky) * sy + gx], bc.y * dex_dy);
                        hz1 = fmaf(c.w, g1, hz1);
                    }
                    if (kz >= 0) {
                        const float2 bc = S.zbc[0][k];
                        const float g0 = fmaf(bc.x, (zk == V.nz ? Q.gh_ghost[2][0] : Q.gh_in[2][0])[(static_cast<size_t>(kz) * V.ny + gy) * sy + gx], bc.y * dex_dz);
                        hy1 = fmaf(-c.w, g0, hy1);
                    }
                }
                hb_w[H_COMP + r * PE + cR] = hy1; hb_w[2 * H_COMP + r * PE + cR] = hz1;
            }
            // this warp holds everything it needs of the old-H slot of z_k in registers: release it to the producer
            __syncwarp();
            if (lane == 0) mbar_arrive(emptyH + (k % LO::NH));
        }

        // ---- phase B: E'(z_k - 1) on the core tile ----------------------------------------------------------------
        if (k >= 1 && warp_core) {
            const int zp = zk - 1;
            // CPML, x slabs: lane j's ψ_E of plane z_k - 1 — requested first, used after the quad loop
            float xp0 = 0.f, xp1 = 0.f;
            if (PML && nxs) {
                if (xlane) {
                    const unsigned qi = static_cast<unsigned>(zp) * xq_plane + xq_row;
                    xp0 = Q.ge[0][0][qi]; xp1 = Q.ge[0][1][qi];
                }
            }
            const float* e_own = reinterpret_cast<const float*>(e_slots + static_cast<size_t>(k % LO::NE) * LO::E_SLOT_BYTES) + (tyb + 1) * PE + 4 + 4 * tx;
            const float* ph = hb + static_cast<size_t>((k - 1) & 1) * 3 * H_COMP + tyb * PE + 4 + 4 * tx;
            float4 ex = ld4(e_own), ey = ld4(e_own + E_COMP), ez = ld4(e_own + 2 * E_COMP);
            // last read of the old-E slot of plane z_k - 1 by this warp (phase A used it as the plane below)
            __syncwarp();
            if (lane == 0) mbar_arrive(emptyE + (k % LO::NE));
            if (BX != 32 && producer && k + 2 < n_h) {
                // two thread rows per warp: the producer warp owns a row of E' too.  It re-arms the rings here — every warp releases
                // these two slots at this point of its own phase B — instead of after its E' arithmetic, half an iteration later
                if (lane == 0) mbar_wait(emptyH + (k % LO::NH), static_cast<uint32_t>((k / LO::NH) & 1));
                issue_h(zk + 2, k % LO::NH);
                if (lane == 0) mbar_wait(emptyE + (k % LO::NE), static_cast<uint32_t>((k / LO::NE) & 1));
                issue_e(zk + 2, k % LO::NE);
            }
            const float4 hz = hzo;
            const float4 hz_yp = ld4(ph + 2 * H_COMP + PE), hx_yp = ld4(ph + PE);
            const float hz_r = ph[2 * H_COMP + 4], hy_r = ph[H_COMP + 4];
            const bool z_in = has_a;                                // H'(z_k) exists inside the lattice
            const bool y_in = y + 1 < V.ny, x_in = x + 4 < V.nx;
            const float idy = y_in ? V.inv_dx[1] : 0.f, idz = z_in ? V.inv_dx[2] : 0.f;
            long long cell = 0;
            int s0 = 0, cnt = 0;
            if (!LEAN) cell = static_cast<long long>(zp) * sz_ll + static_cast<long long>(y) * V.nx + x;
            if (has_src && plane_bit(src_planes, k - 1)) {
                if (LEAN) cell = static_cast<long long>(zp) * sz_ll + static_cast<long long>(y) * V.nx + x;
                s0 = lower_bound_ll(F.cells, F.n, cell);
                while (s0 + cnt < F.n && __ldg(F.cells + s0 + cnt) < cell + 4) ++cnt;
            }
            const float hxv[4] = {hxo.x, hxo.y, hxo.z, hxo.w};
            const float hyv[4] = {hyo.x, hyo.y, hyo.z, hyo.w};
            const float hzv[4] = {hz.x, hz.y, hz.z, hz.w};
            const float hxzp[4] = {hxn.x, hxn.y, hxn.z, hxn.w};
            const float hyzp[4] = {hyn.x, hyn.y, hyn.z, hyn.w};
            const float hzyp[4] = {hz_yp.x, hz_yp.y, hz_yp.z, hz_yp.w};
            const float hxyp[4] = {hx_yp.x, hx_yp.y, hx_yp.z, hx_yp.w};
            const float hzxp[4] = {hz.y, hz.z, hz.w, hz_r};
            const float hyxp[4] = {hyo.y, hyo.z, hyo.w, hy_r};
            float exv[4] = {ex.x, ex.y, ex.z, ex.w};
            float eyv[4] = {ey.x, ey.y, ey.z, ey.w};
            float ezv[4] = {ez.x, ez.y, ez.z, ez.w};
            auto e_quad = [&](auto with_src) {
#pragma unroll
            for (int q = 0; q < 4; ++q) {
                const bool xq_in = XPAD ? (q < xlim) : ((q < 3) || x_in);
                const float dhz_dy = deriv_masked<EXACT>(hzyp[q], hzv[q], V.dx[1], idy, y_in);
                const float dhx_dy = deriv_masked<EXACT>(hxyp[q], hxv[q], V.dx[1], idy, y_in);
                const float dhy_dz = deriv_masked<EXACT>(hyzp[q], hyv[q], V.dx[2], idz, z_in);
                const float dhx_dz = deriv_masked<EXACT>(hxzp[q], hxv[q], V.dx[2], idz, z_in);
                const float dhz_dx = xq_in ? deriv<EXACT>(hzxp[q], hzv[q], V.dx[0], V.inv_dx[0]) : 0.f;
                const float dhy_dx = xq_in ? deriv<EXACT>(hyxp[q], hyv[q], V.dx[0], V.inv_dx[0]) : 0.f;
                const float cx = dhz_dy - dhy_dz;
                const float cy = dhx_dz - dhz_dx;
                const float cz = dhy_dx - dhx_dy;
                float jx = 0.f, jy = 0.f, jz = 0.f;
                if (decltype(with_src)::value && cnt) { const float3 v = forcing_lookup(F, s0, cnt, cell + q, 0); jx = v.x; jy = v.y; jz = v.z; }
                const float4 c = coef(mk_prev[q]);
                exv[q] = leap<EXACT>(c.x, exv[q], c.y, cx, jx);
                eyv[q] = leap<EXACT>(c.x, eyv[q], c.y, cy, jy);
                ezv[q] = leap<EXACT>(c.x, ezv[q], c.y, cz, jz);
            }
            };
            if (!LEAN || has_src) e_quad(std::true_type{}); else e_quad(std::false_type{});
            if (PML && nxs) {
                // x slabs (k_pml_axis<1, 0>): lane j advances ψ_E of the row's j-th x-slab cell in place from ∂H'y/∂x, ∂H'z/∂x of the
                // exchanged H' tile of plane z_k - 1; the cell's owner adds E_y −= c_b·G_xz, E_z += c_b·G_xy
                if (xlane) {
                    const float* pj = hb + static_cast<size_t>((k - 1) & 1) * 3 * H_COMP + ty * PE + 4 + xl_j;
                    const bool in = x0 + xl_j + 1 < nx_cells;
                    const float dhy = in ? deriv<EXACT>(pj[H_COMP + 1], pj[H_COMP], V.dx[0], V.inv_dx[0]) : 0.f;
                    const float dhz = in ? deriv<EXACT>(pj[2 * H_COMP + 1], pj[2 * H_COMP], V.dx[0], V.inv_dx[0]) : 0.f;
                    const float2 bc = S.xbc[1][lane];
                    const float g0 = fmaf(bc.x, xp0, bc.y * dhy);
                    const float g1 = fmaf(bc.x, xp1, bc.y * dhz);
                    S.xg[1][0][ty][lane] = g0; S.xg[1][1][ty][lane] = g1;
                    const unsigned qi = static_cast<unsigned>(zp) * xq_plane + xq_row;
                    Q.ge[0][0][qi] = g0; Q.ge[0][1][qi] = g1;
                }
                __syncwarp();
                if (xjp != -1 && core) {
#pragma unroll
                    for (int q = 0; q < 4; ++q) {
                        const int j = (xjp << (24 - 8 * q)) >> 24;
                        if (j >= 0) {
                            const float w = coef(mk_prev[q]).y;
                            eyv[q] = fmaf(-w, S.xg[1][1][ty][j], eyv[q]);
                            ezv[q] = fmaf(w, S.xg[1][0][ty][j], ezv[q]);
                        }
                    }
                }
            }
            if (PML && (cta_y || zwin_b)) {
                // y and z slabs: ψ_E advanced in place by the owner of E' (k_pml_axis<1, 1>, <1, 2>): E_b −= c_b·G_ac, E_c += c_b·G_ab
                const int pky = core ? S.yk[ty] : -1, pkz = core ? S.zk[k - 1] : -1;
                if (pky >= 0) {
                    const size_t qy = (static_cast<size_t>(zp) * Q.n[1] + pky) * sy + x;
                    const float4 a = ld4(Q.ge[1][0] + qy), b = ld4(Q.ge[1][1] + qy);
                    const float2 bc = S.ybc[1][ty];
                    float g0[4] = {a.x, a.y, a.z, a.w}, g1[4] = {b.x, b.y, b.z, b.w};
#pragma unroll
                    for (int q = 0; q < 4; ++q) {
                        const float w = coef(mk_prev[q]).y;
                        g0[q] = fmaf(bc.x, g0[q], bc.y * deriv_masked<EXACT>(hzyp[q], hzv[q], V.dx[1], idy, y_in));
                        g1[q] = fmaf(bc.x, g1[q], bc.y * deriv_masked<EXACT>(hxyp[q], hxv[q], V.dx[1], idy, y_in));
                        ezv[q] = fmaf(-w, g1[q], ezv[q]);
                        exv[q] = fmaf(w, g0[q], exv[q]);
                    }
                    st4(Q.ge[1][0] + qy, make_float4(g0[0], g0[1], g0[2], g0[3]));
                    st4(Q.ge[1][1] + qy, make_float4(g1[0], g1[1], g1[2], g1[3]));
                }
                if (pkz >= 0) {
                    const size_t qz = (static_cast<size_t>(pkz) * V.ny + y) * sy + x;
                    const float4 a = ld4(Q.ge[2][0] + qz), b = ld4(Q.ge[2][1] + qz);
                    const float2 bc = S.zbc[1][k - 1];
                    float g0[4] = {a.x, a.y, a.z, a.w}, g1[4] = {b.x, b.y, b.z, b.w};
#pragma unroll
                    for (int q = 0; q < 4; ++q) {
                        const float w = coef(mk_prev[q]).y;
                        g0[q] = fmaf(bc.x, g0[q], bc.y * deriv_masked<EXACT>(hxzp[q], hxv[q], V.dx[2], idz, z_in));
                        g1[q] = fmaf(bc.x, g1[q], bc.y * deriv_masked<EXACT>(hyzp[q], hyv[q], V.dx[2], idz, z_in));
                        exv[q] = fmaf(-w, g1[q], exv[q]);
                        eyv[q] = fmaf(w, g0[q], eyv[q]);
                    }
                    st4(Q.ge[2][0] + qz, make_float4(g0[0], g0[1], g0[2], g0[3]));
                    st4(Q.ge[2][1] + qz, make_float4(g1[0], g1[1], g1[2], g1[3]));
                }
            }
            ex = make_float4(exv[0], exv[1], exv[2], exv[3]);
            ey = make_float4(eyv[0], eyv[1], eyv[2], eyv[3]);
            ez = make_float4(ezv[0], ezv[1], ezv[2], ezv[3]);
            if (core) {
                const size_t gi = static_cast<size_t>(zp) * sz + static_cast<size_t>(y) * sy + static_cast<size_t>(x);
                stg4(V.out[0] + gi, ex); stg4(V.out[1] + gi, ey); stg4(V.out[2] + gi, ez);
                if (PUSH) {
                    const size_t pi = static_cast<size_t>(y) * sy + static_cast<size_t>(x);
                    if (zp == 0 && V.push_down[0] != nullptr) { st4(V.push_down[0] + pi, ex); st4(V.push_down[1] + pi, ey); st4(V.push_down[2] + pi, ez); }
                    if (zp == V.nz - 1 && V.push_up[0] != nullptr) { st4(V.push_up[0] + pi, ex); st4(V.push_up[1] + pi, ey); }
                }
                if (has_probe && plane_bit(probe_planes, k - 1)) {
                    if (LEAN) cell = static_cast<long long>(zp) * sz_ll + static_cast<long long>(y) * V.nx + x;
                    for (int s = lower_bound_ll(P.cells, P.n, cell); s < P.n; ++s) {
                        const long long c = __ldg(P.cells + s);
                        if (c >= cell + 4) break;
                        const int q = static_cast<int>(c - cell);
                        float* o = P.row + 6 * static_cast<size_t>(__ldg(P.slot + s));
                        o[0] = pick(ex, q); o[1] = pick(ey, q); o[2] = pick(ez, q);
                        o[3] = pick(hxo, q); o[4] = pick(hyo, q); o[5] = pick(hz, q);
                    }
                }
            }
        }
        else {
            // no phase B in this iteration (k == 0, or the top thread row): this warp is done with the E slot of z_k - 1 too
            __syncwarp();
            if (lane == 0) mbar_arrive(emptyE + (k % LO::NE));
        }
        // Refill as soon as the other warps have drained the slots — the top row's warp has no E' work and waits here in
        // their stead, so the copies of plane z_k + 2 start about half an iteration before the end-of-iteration barrier.
        // H ring slot k % 2 held z_k → z_k + 2;  E ring slot k % 3 held plane z_k - 1 → z_k + 2.
        if (producer && k + 2 < n_h && (BX == 32 || k == 0)) {   // (two rows per warp: done inside phase B, which iteration 0 does not have)
            if (lane == 0) mbar_wait(emptyH + (k % LO::NH), static_cast<uint32_t>((k / LO::NH) & 1));
            issue_h(zk + 2, k % LO::NH);
            if (lane == 0) mbar_wait(emptyE + (k % LO::NE), static_cast<uint32_t>((k / LO::NE) & 1));
            issue_e(zk + 2, k % LO::NE);
        }
        hxo = hxn; hyo = hyn; hzo = hzn;
        mk_prev[0] = mk_cur[0]; mk_prev[1] = mk_cur[1]; mk_prev[2] = mk_cur[2]; mk_prev[3] = mk_cur[3];

        __syncthreads();   // H'(z_k) tile complete before any warp reads it in the next iteration's phase B
    }
#ifdef FDTD_B200_CTA_TRACE
    if (tid == 0 && g_cta_trace) g_cta_trace[trace_at + 2] = globaltimer_ns();
#endif
}

}  // namespace fdtd_b200
