// Fused single-pass step for two-dimensional lattices (nz == 1): the application's own configurations are 2-D
// (cem-app/src/composer/mod.rs:841-869: volume half extent z = 0 → 100×100×1), and a large one (4096×4096×1) is the worst case
// of the z-marching fused kernel — a march of one plane is all ring prologue (67.8 Gcells/s, 0.50 of the HBM roofline).
//
// Here the march runs along y.  A thread owns a quad (4 consecutive x cells) and walks `yc` rows; a WARP covers 32 quads of which
// it owns 31 — the last lane's quad is a halo whose H' it recomputes (from old fields, the step is out of place like the 3-D
// fused kernel) so that lane 30 gets its x+1 neighbour by shuffle; the next warp owns that quad as its lane 0.  Nothing is
// exchanged between warps: no shared memory, no barriers, every lane runs the same instruction stream.  Per row a thread keeps
// H'(y) and old E(y) in registers, loads old E, old H and the index of row y+1 (the loads of row y+2 are issued first), computes
// H'(y+1), then E'(y) from H'(y), H'(y+1) and the shuffled x+1 values, and stores both.  49 B per cell update, as the 3-D kernel.
//
// With nz == 1 every ∂/∂z is zero by the Neumann rule (boundary_condition.rs:29-37: both z neighbours are missing), so
//   H'x = d_a Hx + d_b(−(∂Ez/∂y − 0) − mx)      E'x = c_a Ex + c_b((∂H'z/∂y − 0) − jx)
//   H'y = d_a Hy + d_b(−(0 − ∂Ez/∂x) − my)      E'y = c_a Ey + c_b((0 − ∂H'z/∂x) − jy)
//   H'z = d_a Hz + d_b(−(∂Ey/∂x − ∂Ex/∂y) − mz) E'z = c_a Ez + c_b((∂H'y/∂x − ∂H'x/∂y) − jz)
// evaluated with the same helpers, in the same order, as every other kernel generation (update.wgsl:56-177): bit-identical to
// the fp32 oracle in exact mode and to the other fast-mode kernels in fast mode.
#pragma once

#include "kernels.cuh"
#include "kernels_bulk.cuh"   // mbarrier / bulk-copy helpers

namespace fdtd_b200 {

struct Step2dView {
    const float* in[6];   // Ex Ey Ez Hx Hy Hz of the current set (plane 0)
    float* out[6];        // … of the pending set
    const void* mat;
    const float4* table;
    int n_table;
    int table_in_smem;    // 1: the CTA stages the coefficient table in shared memory (n_table ≤ k2dTableMax)
    int nx, ny;           // nx = row pitch
    int nxl;              // cells of a row that belong to the lattice
    float dx[3], inv_dx[3];
};

constexpr int k2dWarps = 4;          // warps per CTA (independent of each other; 128 threads for occupancy granularity)
constexpr int k2dOwned = 31;         // quads a warp owns (lane 31 is the halo quad)

template <int EXACT>
__device__ __forceinline__ float dsel(bool in, float hi, float lo, float dx, float inv_dx) {
    if (EXACT == 2) return in ? __fdiv_rn(__fsub_rn(hi, lo), dx) : 0.f;
    if (EXACT == 1) return in ? __fmul_rn(__fsub_rn(hi, lo), inv_dx) : 0.f;
    return in ? __fmul_rn(hi - lo, inv_dx) : 0.f;   // uncontracted product, as the select in k_update_*_vec4 leaves it
}

// ---- per-warp shared-memory ring: rows are brought in by 1-D bulk async copies (cp.async.bulk, SASS UBLKCP) ----------------------
// First version of this kernel prefetched the next row into registers: ncu showed it latency-bound (22 % warps active at 128
// registers, 5.7 long-scoreboard stalls per issue, 4.93 TB/s) — registers were both the state and the only prefetch buffer, ≈ 50 KB
// in flight per SM.  Now every warp owns a ring of S row slots in shared memory; lane 0 issues the six copies of a row (512 B each;
// Ey and Ez start one quad early so that the x−1 halo of lane 0 comes with them) completing on the slot's mbarrier, up to S − 1 rows
// ahead of the row being computed: ≈ 6 KB in flight per warp, > 100 KB per SM, none of it in registers.  A slot is released by
// __syncwarp() once every lane has copied its quad out; nothing is shared between warps, so there is still no CTA barrier.
constexpr int k2dStages = 3;
constexpr int k2dSlotFloats = 4 * 128 + 2 * 132;                 // ex hx hy hz: 128 floats; ey ez: 4 halo floats + 128
constexpr int k2dIdxBytes = 288;                                 // the row's material indices: ≤ 15 bytes of alignment slack + 128 × 2 bytes
constexpr int k2dSlotBytes = k2dSlotFloats * 4 + k2dIdxBytes;
constexpr int k2dWarpBytes = k2dStages * k2dSlotBytes + 64;      // + S mbarriers (padded)
constexpr int k2dTableMax = 512;                                 // coefficient rows staged in shared memory (8 KB)
__host__ __device__ constexpr int k2d_smem_bytes(int n_table_smem) { return n_table_smem * 16 + k2dWarps * k2dWarpBytes; }

template <typename IdxT> struct IdxQuad;
template <> struct IdxQuad<uint8_t> {
    uint32_t raw = 0;
    __device__ __forceinline__ void load(const void* base, size_t i) { raw = *reinterpret_cast<const uint32_t*>(static_cast<const uint8_t*>(base) + i); }
    __device__ __forceinline__ int get(int k) const { return static_cast<int>((raw >> (8 * k)) & 0xffu); }
};
template <> struct IdxQuad<uint16_t> {
    uint2 raw = {0u, 0u};
    __device__ __forceinline__ void load(const void* base, size_t i) { raw = *reinterpret_cast<const uint2*>(static_cast<const uint16_t*>(base) + i); }
    __device__ __forceinline__ int get(int k) const { return static_cast<int>(((k < 2 ? raw.x : raw.y) >> (16 * (k & 1))) & 0xffffu); }
};

__device__ __forceinline__ void st4cs(float* p, const float4& v) { __stcs(reinterpret_cast<float4*>(p), v); }

template <int EXACT, typename IdxT>
__global__ void __launch_bounds__(32 * k2dWarps, 4)
k_step_2d(const Step2dView V, const ForcingView F, const ProbeView P, const int yc) {
    griddep_launch_dependents();
    extern __shared__ __align__(128) unsigned char smem_2d[];
    const unsigned lane = threadIdx.x & 31u;
    const int wic = threadIdx.x >> 5;
    const float4* __restrict__ tab = V.table;
    const int tab_bytes = V.table_in_smem ? V.n_table * 16 : 0;
    if (V.table_in_smem) {                                       // the only CTA-wide step: before any warp may leave
        float4* s_tab = reinterpret_cast<float4*>(smem_2d);
        for (int t = threadIdx.x; t < V.n_table; t += 32 * k2dWarps) s_tab[t] = V.table[t];
        __syncthreads();
        tab = s_tab;
    }
    const int nq = V.nx / 4;
    // work items = (x segment of 31 quads, chunk of yc rows), dealt to warps linearly (x fastest): the warps of a CTA are independent,
    // so a lattice whose width is not a multiple of 4 segments wastes no warp slots
    const int n_xseg = (nq + k2dOwned - 1) / k2dOwned;
    const int item = blockIdx.x * k2dWarps + wic;
    const int warp = item % n_xseg;                              // x segment
    const int q = warp * k2dOwned + static_cast<int>(lane);      // quad index along x
    const int ya = (item / n_xseg) * yc;
    const int yb = min(ya + yc, V.ny);
    if (ya >= yb) return;                                        // warp-uniform; warps never synchronise with each other
    const bool act = q < nq;                                     // surplus lanes shadow the last quad and never store
    const int x = (act ? q : nq - 1) * 4;
    const bool own = act && lane < k2dOwned;                     // lane 31: halo quad, H' only, no stores
    const size_t sy = static_cast<size_t>(V.nx);
    const int x_lo = warp * k2dOwned * 4;
    const bool x_in = x > 0;
    const bool halo_x = x_lo > 0;                                // the warp's rows start one quad early for Ey, Ez

    unsigned char* wbase = smem_2d + tab_bytes + static_cast<size_t>(wic) * k2dWarpBytes;
    uint64_t* full = reinterpret_cast<uint64_t*>(wbase + k2dStages * k2dSlotBytes);
    // the index row of the segment starts at byte x_lo·sizeof(IdxT) of its lattice row, 4- or 8-byte aligned only: the bulk copy
    // starts at the 16-byte boundary below it (rows of the index are multiples of 16 bytes on fused handles)
    const uint32_t idx_off = static_cast<uint32_t>(static_cast<size_t>(x_lo) * sizeof(IdxT)) & 15u;
    const int n_rows = min(yb + 1, V.ny) - ya;                   // rows ya … yb (the row above the chunk feeds the recomputed H')
    const uint32_t row_bytes = static_cast<uint32_t>(min(128, V.nx - x_lo)) * 4u;
    const int lq = (act ? q : nq - 1) - warp * k2dOwned;         // this lane's quad inside the warp's row segment

    const uint32_t idx_bytes = (idx_off + (row_bytes / 4u) * static_cast<uint32_t>(sizeof(IdxT)) + 15u) & ~15u;
    auto issue = [&](int t) {                                    // lane 0: request row ya + t into slot t % S
        float* sl = reinterpret_cast<float*>(wbase + (t % k2dStages) * k2dSlotBytes);
        uint64_t* bar = full + (t % k2dStages);
        const size_t g = static_cast<size_t>(ya + t) * sy + static_cast<size_t>(x_lo);
        const uint32_t hb = row_bytes + (halo_x ? 16u : 0u);
        mbar_arrive_expect_tx(bar, 4u * row_bytes + 2u * hb + idx_bytes);
        bulk_g2s(sl + k2dSlotFloats, static_cast<const unsigned char*>(V.mat) + g * sizeof(IdxT) - idx_off, idx_bytes, bar);
        bulk_g2s(sl, V.in[0] + g, row_bytes, bar);
        bulk_g2s(sl + 128 + (halo_x ? 0 : 4), V.in[1] + g - (halo_x ? 4 : 0), hb, bar);
        bulk_g2s(sl + 260 + (halo_x ? 0 : 4), V.in[2] + g - (halo_x ? 4 : 0), hb, bar);
        bulk_g2s(sl + 392, V.in[3] + g, row_bytes, bar);
        bulk_g2s(sl + 520, V.in[4] + g, row_bytes, bar);
        bulk_g2s(sl + 648, V.in[5] + g, row_bytes, bar);
    };
    griddep_wait();   // the previous step's grid is complete (kernels.cuh): everything above overlapped its tail
    if (lane == 0) {
        for (int k = 0; k < k2dStages; ++k) mbar_init(full + k, 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
        for (int t = 0; t < min(k2dStages, n_rows); ++t) issue(t);
    }
    __syncwarp();

    // forcing / probes: one scan per warp over the (short, sorted) lists, so that rows without any cost nothing
    bool has_src = false, has_probe = false;
    const long long plane_ll = static_cast<long long>(V.nx) * V.ny;
    if (F.n) has_src = __any_sync(0xffffffffu, any_cell_in_box(F.cells, F.n, lane, 32, x_lo, x_lo + 128, ya, yb + 1, 0, 1, V.nx, plane_ll));
    if (P.n) has_probe = __any_sync(0xffffffffu, any_cell_in_box(P.cells, P.n, lane, 32, x_lo, x_lo + 128, ya, yb, 0, 1, V.nx, plane_ll));

    struct RowE { float4 ex, ey, ez; };
    struct RowH { float4 hx, hy, hz; };
    // wait for row t, copy this lane's quad (and its x−1 halo) out of the slot, release the slot and refill it with row t + S
    auto take = [&](int t, RowE& e, RowH& h, float& ey_l, float& ez_l, IdxQuad<IdxT>& mk) {
        const float* sl = reinterpret_cast<const float*>(wbase + (t % k2dStages) * k2dSlotBytes);
        mbar_wait(full + (t % k2dStages), static_cast<uint32_t>((t / k2dStages) & 1));
        e.ex = ld4(sl + 4 * lq); e.ey = ld4(sl + 132 + 4 * lq); e.ez = ld4(sl + 264 + 4 * lq);
        h.hx = ld4(sl + 392 + 4 * lq); h.hy = ld4(sl + 520 + 4 * lq); h.hz = ld4(sl + 648 + 4 * lq);
        ey_l = sl[132 + 4 * lq - 1]; ez_l = sl[264 + 4 * lq - 1];    // x−1 of the quad: the previous quad's .w, or the halo float
        mk.load(reinterpret_cast<const unsigned char*>(sl + k2dSlotFloats) + idx_off, static_cast<size_t>(4 * lq));
        __syncwarp();
        if (lane == 0 && t + k2dStages < n_rows) issue(t + k2dStages);
    };

    // H' of one row from its old fields, the old E of the row below it (zero vector where y == 0)
    auto h_row = [&](const RowE& e, const RowH& ho, float ey_l, float ez_l, const IdxQuad<IdxT>& mk, const float4& ex_ym, const float4& ez_ym, bool y_in,
                     size_t i, RowH& hn) {
        const float exv[4] = {e.ex.x, e.ex.y, e.ex.z, e.ex.w}, eyv[4] = {e.ey.x, e.ey.y, e.ey.z, e.ey.w}, ezv[4] = {e.ez.x, e.ez.y, e.ez.z, e.ez.w};
        const float exym[4] = {ex_ym.x, ex_ym.y, ex_ym.z, ex_ym.w}, ezym[4] = {ez_ym.x, ez_ym.y, ez_ym.z, ez_ym.w};
        const float eyxm[4] = {ey_l, e.ey.x, e.ey.y, e.ey.z}, ezxm[4] = {ez_l, e.ez.x, e.ez.y, e.ez.z};
        float hxv[4] = {ho.hx.x, ho.hx.y, ho.hx.z, ho.hx.w}, hyv[4] = {ho.hy.x, ho.hy.y, ho.hy.z, ho.hy.w}, hzv[4] = {ho.hz.x, ho.hz.y, ho.hz.z, ho.hz.w};
        int s0 = 0, cnt = 0;
        if (has_src) {
            s0 = lower_bound_ll(F.cells, F.n, static_cast<long long>(i));
            while (s0 + cnt < F.n && __ldg(F.cells + s0 + cnt) < static_cast<long long>(i) + 4) ++cnt;
        }
#pragma unroll
        for (int k = 0; k < 4; ++k) {
            const bool xk_in = (k > 0) || x_in;
            const float dez_dy = dsel<EXACT>(y_in, ezv[k], ezym[k], V.dx[1], V.inv_dx[1]);
            const float dex_dy = dsel<EXACT>(y_in, exv[k], exym[k], V.dx[1], V.inv_dx[1]);
            const float dez_dx = dsel<EXACT>(xk_in, ezv[k], ezxm[k], V.dx[0], V.inv_dx[0]);
            const float dey_dx = dsel<EXACT>(xk_in, eyv[k], eyxm[k], V.dx[0], V.inv_dx[0]);
            const float cx = dez_dy - 0.f;          // ∂Ey/∂z = 0
            const float cy = 0.f - dez_dx;          // ∂Ex/∂z = 0
            const float cz = dey_dx - dex_dy;
            float mx = 0.f, my = 0.f, mz = 0.f;
            if (cnt) { const float3 v = forcing_lookup(F, s0, cnt, static_cast<long long>(i) + k, 3); mx = v.x; my = v.y; mz = v.z; }
            const float4 c = tab[mk.get(k)];
            hxv[k] = leap<EXACT>(c.z, hxv[k], c.w, -cx, mx);
            hyv[k] = leap<EXACT>(c.z, hyv[k], c.w, -cy, my);
            hzv[k] = leap<EXACT>(c.z, hzv[k], c.w, -cz, mz);
        }
        hn.hx = make_float4(hxv[0], hxv[1], hxv[2], hxv[3]);
        hn.hy = make_float4(hyv[0], hyv[1], hyv[2], hyv[3]);
        hn.hz = make_float4(hzv[0], hzv[1], hzv[2], hzv[3]);
    };

    // ---- prologue: old E of row ya−1 (backward y-difference of the first H' row), the indices of rows ya and ya+1, H'(ya)
    size_t i = static_cast<size_t>(ya) * sy + static_cast<size_t>(x);
    float4 ex_below = make_float4(0.f, 0.f, 0.f, 0.f), ez_below = ex_below;
    if (ya > 0) { ex_below = ld4(V.in[0] + i - sy); ez_below = ld4(V.in[2] + i - sy); }
    IdxQuad<IdxT> mk_cur, mk_nxt;
    RowE cur; RowH hold, h0;
    float ey_l, ez_l;
    take(0, cur, hold, ey_l, ez_l, mk_cur);
    h_row(cur, hold, ey_l, ez_l, mk_cur, ex_below, ez_below, ya > 0, i, h0);
    if (own) { st4cs(V.out[3] + i, h0.hx); st4cs(V.out[4] + i, h0.hy); st4cs(V.out[5] + i, h0.hz); }

    for (int y = ya; y < yb; ++y, i += sy) {
        const bool y1_in = y + 1 < V.ny;
        RowE nxt; RowH h1;
        nxt.ex = nxt.ey = nxt.ez = make_float4(0.f, 0.f, 0.f, 0.f);
        h1.hx = h1.hy = h1.hz = make_float4(0.f, 0.f, 0.f, 0.f);
        if (y1_in) {
            take(y + 1 - ya, nxt, hold, ey_l, ez_l, mk_nxt);
            h_row(nxt, hold, ey_l, ez_l, mk_nxt, cur.ex, cur.ez, true, i + sy, h1);
            if (own && y + 1 < yb) { st4cs(V.out[3] + i + sy, h1.hx); st4cs(V.out[4] + i + sy, h1.hy); st4cs(V.out[5] + i + sy, h1.hz); }   // row yb belongs to the next chunk
        }
        // E'(y): forward differences on the new H — x+1 from the next lane's quad, y+1 from the row just computed
        const float hy_r = __shfl_down_sync(0xffffffffu, h0.hy.x, 1);
        const float hz_r = __shfl_down_sync(0xffffffffu, h0.hz.x, 1);
        const float hxv[4] = {h0.hx.x, h0.hx.y, h0.hx.z, h0.hx.w}, hyv[4] = {h0.hy.x, h0.hy.y, h0.hy.z, h0.hy.w}, hzv[4] = {h0.hz.x, h0.hz.y, h0.hz.z, h0.hz.w};
        const float hxyp[4] = {h1.hx.x, h1.hx.y, h1.hx.z, h1.hx.w}, hzyp[4] = {h1.hz.x, h1.hz.y, h1.hz.z, h1.hz.w};
        const float hyxp[4] = {h0.hy.y, h0.hy.z, h0.hy.w, hy_r}, hzxp[4] = {h0.hz.y, h0.hz.z, h0.hz.w, hz_r};
        float exv[4] = {cur.ex.x, cur.ex.y, cur.ex.z, cur.ex.w}, eyv[4] = {cur.ey.x, cur.ey.y, cur.ey.z, cur.ey.w}, ezv[4] = {cur.ez.x, cur.ez.y, cur.ez.z, cur.ez.w};
        int s0 = 0, cnt = 0;
        if (has_src) {
            s0 = lower_bound_ll(F.cells, F.n, static_cast<long long>(i));
            while (s0 + cnt < F.n && __ldg(F.cells + s0 + cnt) < static_cast<long long>(i) + 4) ++cnt;
        }
#pragma unroll
        for (int k = 0; k < 4; ++k) {
            const bool xk_in = x + k + 1 < V.nxl;
            const float dhz_dy = dsel<EXACT>(y1_in, hzyp[k], hzv[k], V.dx[1], V.inv_dx[1]);
            const float dhx_dy = dsel<EXACT>(y1_in, hxyp[k], hxv[k], V.dx[1], V.inv_dx[1]);
            const float dhz_dx = dsel<EXACT>(xk_in, hzxp[k], hzv[k], V.dx[0], V.inv_dx[0]);
            const float dhy_dx = dsel<EXACT>(xk_in, hyxp[k], hyv[k], V.dx[0], V.inv_dx[0]);
            const float cx = dhz_dy - 0.f;
            const float cy = 0.f - dhz_dx;
            const float cz = dhy_dx - dhx_dy;
            float jx = 0.f, jy = 0.f, jz = 0.f;
            if (cnt) { const float3 v = forcing_lookup(F, s0, cnt, static_cast<long long>(i) + k, 0); jx = v.x; jy = v.y; jz = v.z; }
            const float4 c = tab[mk_cur.get(k)];
            exv[k] = leap<EXACT>(c.x, exv[k], c.y, cx, jx);
            eyv[k] = leap<EXACT>(c.x, eyv[k], c.y, cy, jy);
            ezv[k] = leap<EXACT>(c.x, ezv[k], c.y, cz, jz);
        }
        if (own) {
            const float4 ex4 = make_float4(exv[0], exv[1], exv[2], exv[3]), ey4 = make_float4(eyv[0], eyv[1], eyv[2], eyv[3]),
                         ez4 = make_float4(ezv[0], ezv[1], ezv[2], ezv[3]);
            st4cs(V.out[0] + i, ex4); st4cs(V.out[1] + i, ey4); st4cs(V.out[2] + i, ez4);
            if (has_probe) {
                for (int s = lower_bound_ll(P.cells, P.n, static_cast<long long>(i)); s < P.n && __ldg(P.cells + s) < static_cast<long long>(i) + 4; ++s) {
                    const int k = static_cast<int>(__ldg(P.cells + s) - static_cast<long long>(i));   // selects, not indexed arrays: no stack frame
                    float* o = P.row + 6 * static_cast<size_t>(__ldg(P.slot + s));
                    o[0] = pick(ex4, k); o[1] = pick(ey4, k); o[2] = pick(ez4, k); o[3] = pick(h0.hx, k); o[4] = pick(h0.hy, k); o[5] = pick(h0.hz, k);
                }
            }
        }
        cur = nxt; h0 = h1;
        mk_cur = mk_nxt;
    }
}

}  // namespace fdtd_b200
