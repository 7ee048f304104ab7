// Device code of fdtd_b200: the Yee leapfrog half-steps (sm_100a), field gather/scatter and
// projection.  Semantics follow cem-solver/src/fdtd/cpu/mod.rs:416-515 and
// cem-solver/src/fdtd/wgpu/update.wgsl:56-177; the design (SoA fp32, in-place update, u8/u16
// material index into a shared-memory coefficient table, z-streaming with the z∓1 plane kept in
// registers, warp shuffles for x neighbours, forcing and probes fused) is described in DESIGN.md.
#pragma once

#include <cuda_runtime.h>
#include <stdint.h>

namespace fdtd_b200 {

// What a half-step kernel needs to know about the slab it works on.
struct LatticeView {
    float* ex; float* ey; float* ez;  // pointers to local plane 0 (ghost planes live at -1 and nz)
    float* hx; float* hy; float* hz;
    const void* mat;      // u8 / u16 material index of local plane 0
    const float4* table;  // (c_a, c_b, d_a, d_b) per material, fp32
    int n_table;
    int table_in_smem;    // 1: kernels stage the table in shared memory (n_table small enough)
    int nx, ny, nz;       // local dims (nz = owned planes); nx is the row pitch of every array
    int nxl;              // cells of a row that belong to the lattice (nxl ≤ nx; cells nxl..nx-1 are pad cells whose material is
                          // the all-zero coefficient row, so they stay 0; only the x+1 Neumann test needs to know them)
    int lo_open, hi_open; // a neighbouring slab exists below / above (ghost plane is valid)
    float dx[3];
    float inv_dx[3];
};

// Per-step forcing: `cells` sorted ascending, unique, local linear indices; vals[k] = (j xyz, m xyz).
struct ForcingView {
    int n;
    const long long* cells;
    const float* vals;
};

// Probes: `cells` sorted ascending (local linear indices), slot[k] = position in the caller's order.
struct ProbeView {
    int n;
    const long long* cells;
    const int* slot;
    float* row;  // ring row of the step being computed: [n_probes_total][6]
};

__device__ __forceinline__ float4 ld4(const float* p) { return *reinterpret_cast<const float4*>(p); }
__device__ __forceinline__ void st4(float* p, const float4& v) { *reinterpret_cast<float4*>(p) = v; }

__device__ __forceinline__ int lower_bound_ll(const long long* __restrict__ a, int n, long long key) {
    int lo = 0, hi = n;
    while (lo < hi) {
        const int mid = (lo + hi) >> 1;
        if (__ldg(a + mid) < key) lo = mid + 1; else hi = mid;
    }
    return lo;
}

// Arithmetic modes (template parameter EXACT):
//   0  fast:  (hi - lo) * (1/dx), FMA contraction allowed;
//   1  exact, every dx a power of two: x * (1/dx) is bit-identical to x / dx, so multiply, no contraction;
//   2  exact, general dx: IEEE division as update.wgsl:159,171.
// (f_hi - f_lo) / dx:
template <int EXACT>
__device__ __forceinline__ float deriv(float hi, float lo, float dx, float inv_dx) {
    if (EXACT == 2) return __fdiv_rn(__fsub_rn(hi, lo), dx);
    if (EXACT == 1) return __fmul_rn(__fsub_rn(hi, lo), inv_dx);
    return (hi - lo) * inv_dx;
}

// The same derivative where the Neumann rule is folded into the reciprocal: `inv_masked` is 1/dx, or 0 where the neighbour does
// not exist (the product is then ±0 instead of a selected +0 — equal as numbers; no contraction, so the non-zero case rounds
// exactly like deriv<>).  EXACT == 2 divides and needs the select.
template <int EXACT>
__device__ __forceinline__ float deriv_masked(float hi, float lo, float dx, float inv_masked, bool in) {
    if (EXACT == 2) return in ? __fdiv_rn(__fsub_rn(hi, lo), dx) : 0.f;
    if (EXACT == 1) return __fmul_rn(__fsub_rn(hi, lo), inv_masked);
    return __fmul_rn(hi - lo, inv_masked);
}

// a*prev + b*(c - s + psi), psi = 0 (update.wgsl:89,127; cpu/mod.rs:460-461, 502-503).
template <int EXACT>
__device__ __forceinline__ float leap(float a, float prev, float b, float c, float s) {
    if (EXACT) {
        float t = __fsub_rn(c, s);
        t = __fadd_rn(t, 0.0f);
        return __fadd_rn(__fmul_rn(a, prev), __fmul_rn(b, t));
    }
    return fmaf(a, prev, b * (c - s));
}

template <typename IdxT> struct Idx4;
template <> struct Idx4<uint8_t> {
    static __device__ __forceinline__ void load(const void* base, size_t i, int (&k)[4]) {
        const uchar4 v = *reinterpret_cast<const uchar4*>(static_cast<const uint8_t*>(base) + i);
        k[0] = v.x; k[1] = v.y; k[2] = v.z; k[3] = v.w;
    }
};
template <> struct Idx4<uint16_t> {
    static __device__ __forceinline__ void load(const void* base, size_t i, int (&k)[4]) {
        const ushort4 v = *reinterpret_cast<const ushort4*>(static_cast<const uint16_t*>(base) + i);
        k[0] = v.x; k[1] = v.y; k[2] = v.z; k[3] = v.w;
    }
};

// Value (offset 0 → j, 3 → m) of the forcing entry registered for `cell`, zero if none.
__device__ __noinline__ float3 forcing_lookup(const ForcingView F, int s0, int cnt, long long cell, int off) {
    float3 v = make_float3(0.f, 0.f, 0.f);
    for (int s = s0; s < s0 + cnt; ++s) {
        if (__ldg(F.cells + s) == cell) {
            v.x = __ldg(F.vals + 6 * s + off);
            v.y = __ldg(F.vals + 6 * s + off + 1);
            v.z = __ldg(F.vals + 6 * s + off + 2);
        }
    }
    return v;
}

// Does any of the (sorted) cells lie inside the CTA's box [x0,x1) x [y0,y1) x [z0,z1)?  Evaluated once per CTA so
// that the per-plane forcing / probe lookups cost nothing in the (vast majority of) CTAs that own none.
__device__ __forceinline__ bool any_cell_in_box(const long long* __restrict__ cells, int n, int tid, int nthreads, int x0, int x1,
                                                int y0, int y1, int z0, int z1, int nx, long long sz) {
    bool hit = false;
    for (int s = tid; s < n; s += nthreads) {
        const long long c = __ldg(cells + s);
        const int z = static_cast<int>(c / sz);
        const long long r = c - static_cast<long long>(z) * sz;
        const int yy = static_cast<int>(r / nx);
        const int xx = static_cast<int>(r - static_cast<long long>(yy) * nx);
        hit |= (z >= z0) && (z < z1) && (yy >= y0) && (yy < y1) && (xx >= x0) && (xx < x1);
    }
    return hit;
}

// Programmatic dependent launch (griddepcontrol): a step kernel lets the next launch of the stream become resident as soon as all of
// its own CTAs have started (launch_dependents at entry), and waits for the previous grid to complete and flush (wait) before its
// first access to anything a step writes — fields, ψ, probe rows.  What runs before the wait (barrier set-up, coefficient table,
// the per-CTA scans of the constant source / probe cell lists) overlaps the previous step's tail and the launch latency.
__device__ __forceinline__ void griddep_launch_dependents() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }
__device__ __forceinline__ void griddep_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }

// The same scan, also recording WHICH planes of the box hold a cell: bit (z − z0) of `planes` (shared memory, zeroed by the
// caller, ≥ (z1 − z0 + 31) / 32 words).  A marching kernel then pays for the per-cell lookups on those planes only.
__device__ __forceinline__ bool mark_planes_in_box(const long long* __restrict__ cells, int n, int tid, int nthreads, int x0, int x1,
                                                   int y0, int y1, int z0, int z1, int nx, long long sz, unsigned* planes) {
    bool hit = false;
    for (int s = tid; s < n; s += nthreads) {
        const long long c = __ldg(cells + s);
        const int z = static_cast<int>(c / sz);
        const long long r = c - static_cast<long long>(z) * sz;
        const int yy = static_cast<int>(r / nx);
        const int xx = static_cast<int>(r - static_cast<long long>(yy) * nx);
        if ((z >= z0) && (z < z1) && (yy >= y0) && (yy < y1) && (xx >= x0) && (xx < x1)) {
            hit = true;
            atomicOr(planes + ((z - z0) >> 5), 1u << ((z - z0) & 31));
        }
    }
    return hit;
}

__device__ __forceinline__ float pick(const float4& v, int k) { return k == 0 ? v.x : (k == 1 ? v.y : (k == 2 ? v.z : v.w)); }

// ------------------------------------------------------------------------------------------------
// float4 z-streaming kernels (nx % 4 == 0).  Thread = 4 consecutive x cells of one (y) row, marching
// over `zc` planes.  BX % 32 == 0 so a warp covers 128 consecutive x cells of one row.
// ------------------------------------------------------------------------------------------------

// H' = d_a·H + d_b·(−∇×E − M); backward differences, zero where x_i == 0 (update.wgsl:56-91, 155-165).
template <int EXACT, typename IdxT, int BX, int BY>
__global__ void __launch_bounds__(BX * BY)
k_update_h_vec4(const LatticeView L, const ForcingView F, const int z_lo, const int z_hi, const int zc) {
    extern __shared__ float2 s_tab[];
    if (L.table_in_smem) {
        for (int t = threadIdx.y * BX + threadIdx.x; t < L.n_table; t += BX * BY) {
            const float4 c = L.table[t];
            s_tab[t] = make_float2(c.z, c.w);
        }
    }
    __syncthreads();
    const int za = z_lo + blockIdx.z * zc;
    const int zb = min(za + zc, z_hi);
    const long long sz_ll = static_cast<long long>(L.nx) * L.ny;
    bool has_src = false;
    if (F.n) has_src = __syncthreads_or(any_cell_in_box(F.cells, F.n, threadIdx.y * BX + threadIdx.x, BX * BY, blockIdx.x * BX * 4,
                                                        (blockIdx.x + 1) * BX * 4, blockIdx.y * BY, (blockIdx.y + 1) * BY, za, zb, L.nx, sz_ll)) != 0;

    const int y = blockIdx.y * BY + threadIdx.y;
    if (y >= L.ny) return;
    const int x_raw = (blockIdx.x * BX + threadIdx.x) * 4;
    const bool act = x_raw < L.nx;
    const int x = act ? x_raw : L.nx - 4;  // surplus lanes shadow the last quad and never store
    if (za >= zb) return;
    const unsigned lane = threadIdx.x & 31u;
    const size_t sy = static_cast<size_t>(L.nx);
    const size_t sz = sy * static_cast<size_t>(L.ny);
    size_t i = static_cast<size_t>(za) * sz + static_cast<size_t>(y) * sy + static_cast<size_t>(x);

    const bool y_in = y > 0;
    const bool x_in = x > 0;
    bool z_in = (za > 0) || (L.lo_open != 0);
    float4 ex_zm = make_float4(0.f, 0.f, 0.f, 0.f), ey_zm = ex_zm;
    if (z_in) { ex_zm = ld4(L.ex + i - sz); ey_zm = ld4(L.ey + i - sz); }

    for (int z = za; z < zb; ++z, i += sz) {
        const float4 ex = ld4(L.ex + i), ey = ld4(L.ey + i), ez = ld4(L.ez + i);
        float4 hx = ld4(L.hx + i), hy = ld4(L.hy + i), hz = ld4(L.hz + i);
        int mk[4];
        Idx4<IdxT>::load(L.mat, i, mk);
        float4 ex_ym = ex, ez_ym = ez;
        if (y_in) { ex_ym = ld4(L.ex + i - sy); ez_ym = ld4(L.ez + i - sy); }
        float ey_l = __shfl_up_sync(0xffffffffu, ey.w, 1);
        float ez_l = __shfl_up_sync(0xffffffffu, ez.w, 1);
        if (lane == 0 && x_in) { ey_l = L.ey[i - 1]; ez_l = L.ez[i - 1]; }

        int s0 = 0, cnt = 0;
        if (has_src) {
            s0 = lower_bound_ll(F.cells, F.n, static_cast<long long>(i));
            while (s0 + cnt < F.n && __ldg(F.cells + s0 + cnt) < static_cast<long long>(i) + 4) ++cnt;
        }

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
#pragma unroll
        for (int k = 0; k < 4; ++k) {
            const bool xk_in = (k > 0) || x_in;
            const float dez_dy = y_in ? deriv<EXACT>(ezv[k], ezym[k], L.dx[1], L.inv_dx[1]) : 0.f;
            const float dex_dy = y_in ? deriv<EXACT>(exv[k], exym[k], L.dx[1], L.inv_dx[1]) : 0.f;
            const float dey_dz = z_in ? deriv<EXACT>(eyv[k], eyzm[k], L.dx[2], L.inv_dx[2]) : 0.f;
            const float dex_dz = z_in ? deriv<EXACT>(exv[k], exzm[k], L.dx[2], L.inv_dx[2]) : 0.f;
            const float dez_dx = xk_in ? deriv<EXACT>(ezv[k], ezxm[k], L.dx[0], L.inv_dx[0]) : 0.f;
            const float dey_dx = xk_in ? deriv<EXACT>(eyv[k], eyxm[k], L.dx[0], L.inv_dx[0]) : 0.f;
            const float cx = dez_dy - dey_dz;
            const float cy = dex_dz - dez_dx;
            const float cz = dey_dx - dex_dy;
            float mx = 0.f, my = 0.f, mz = 0.f;
            if (cnt) { const float3 v = forcing_lookup(F, s0, cnt, static_cast<long long>(i) + k, 3); mx = v.x; my = v.y; mz = v.z; }
            float2 c;
            if (L.table_in_smem) c = s_tab[mk[k]];
            else { const float4 t4 = __ldg(L.table + mk[k]); c = make_float2(t4.z, t4.w); }
            hxv[k] = leap<EXACT>(c.x, hxv[k], c.y, -cx, mx);
            hyv[k] = leap<EXACT>(c.x, hyv[k], c.y, -cy, my);
            hzv[k] = leap<EXACT>(c.x, hzv[k], c.y, -cz, mz);
        }
        if (act) {
            st4(L.hx + i, make_float4(hxv[0], hxv[1], hxv[2], hxv[3]));
            st4(L.hy + i, make_float4(hyv[0], hyv[1], hyv[2], hyv[3]));
            st4(L.hz + i, make_float4(hzv[0], hzv[1], hzv[2], hzv[3]));
        }
        ex_zm = ex; ey_zm = ey; z_in = true;
    }
}

// E' = c_a·E + c_b·(∇×H' − J); forward differences on the new H, zero where x_i + 1 == n_i
// (update.wgsl:94-129, 167-177).  Probe sampling of (E', H') is fused here.
template <int EXACT, typename IdxT, int BX, int BY>
__global__ void __launch_bounds__(BX * BY)
k_update_e_vec4(const LatticeView L, const ForcingView F, const ProbeView P, const int z_lo, const int z_hi, const int zc) {
    extern __shared__ float2 s_tab[];
    if (L.table_in_smem) {
        for (int t = threadIdx.y * BX + threadIdx.x; t < L.n_table; t += BX * BY) {
            const float4 c = L.table[t];
            s_tab[t] = make_float2(c.x, c.y);
        }
    }
    __syncthreads();
    const int za = z_lo + blockIdx.z * zc;
    const int zb = min(za + zc, z_hi);
    const long long sz_ll = static_cast<long long>(L.nx) * L.ny;
    bool has_src = false, has_probe = false;
    if (F.n) has_src = __syncthreads_or(any_cell_in_box(F.cells, F.n, threadIdx.y * BX + threadIdx.x, BX * BY, blockIdx.x * BX * 4,
                                                        (blockIdx.x + 1) * BX * 4, blockIdx.y * BY, (blockIdx.y + 1) * BY, za, zb, L.nx, sz_ll)) != 0;
    if (P.n) has_probe = __syncthreads_or(any_cell_in_box(P.cells, P.n, threadIdx.y * BX + threadIdx.x, BX * BY, blockIdx.x * BX * 4,
                                                          (blockIdx.x + 1) * BX * 4, blockIdx.y * BY, (blockIdx.y + 1) * BY, za, zb, L.nx, sz_ll)) != 0;

    const int y = blockIdx.y * BY + threadIdx.y;
    if (y >= L.ny) return;
    const int x_raw = (blockIdx.x * BX + threadIdx.x) * 4;
    const bool act = x_raw < L.nx;
    const int x = act ? x_raw : L.nx - 4;
    if (za >= zb) return;
    const unsigned lane = threadIdx.x & 31u;
    const size_t sy = static_cast<size_t>(L.nx);
    const size_t sz = sy * static_cast<size_t>(L.ny);
    size_t i = static_cast<size_t>(za) * sz + static_cast<size_t>(y) * sy + static_cast<size_t>(x);

    const bool y_in = y + 1 < L.ny;
    const bool x_in = x + 4 < L.nx;  // a float4 to the right of the quad exists
    const int xlim = L.nxl - 1 - x;  // cell k of the quad has an x+1 neighbour inside the lattice iff k < xlim
    float4 hx_c = ld4(L.hx + i), hy_c = ld4(L.hy + i);

    for (int z = za; z < zb; ++z, i += sz) {
        const bool z_in = (z + 1 < L.nz) || (L.hi_open != 0);
        float4 hx_n = hx_c, hy_n = hy_c;
        if (z_in) { hx_n = ld4(L.hx + i + sz); hy_n = ld4(L.hy + i + sz); }
        const float4 hz = ld4(L.hz + i);
        float4 ex = ld4(L.ex + i), ey = ld4(L.ey + i), ez = ld4(L.ez + i);
        int mk[4];
        Idx4<IdxT>::load(L.mat, i, mk);
        float4 hz_yp = hz, hx_yp = hx_c;
        if (y_in) { hz_yp = ld4(L.hz + i + sy); hx_yp = ld4(L.hx + i + sy); }
        float hz_r = __shfl_down_sync(0xffffffffu, hz.x, 1);
        float hy_r = __shfl_down_sync(0xffffffffu, hy_c.x, 1);
        if (lane == 31 && x_in) { hz_r = L.hz[i + 4]; hy_r = L.hy[i + 4]; }

        int s0 = 0, cnt = 0;
        if (has_src) {
            s0 = lower_bound_ll(F.cells, F.n, static_cast<long long>(i));
            while (s0 + cnt < F.n && __ldg(F.cells + s0 + cnt) < static_cast<long long>(i) + 4) ++cnt;
        }

        const float hxv[4] = {hx_c.x, hx_c.y, hx_c.z, hx_c.w};
        const float hyv[4] = {hy_c.x, hy_c.y, hy_c.z, hy_c.w};
        const float hzv[4] = {hz.x, hz.y, hz.z, hz.w};
        const float hxzp[4] = {hx_n.x, hx_n.y, hx_n.z, hx_n.w};
        const float hyzp[4] = {hy_n.x, hy_n.y, hy_n.z, hy_n.w};
        const float hzyp[4] = {hz_yp.x, hz_yp.y, hz_yp.z, hz_yp.w};
        const float hxyp[4] = {hx_yp.x, hx_yp.y, hx_yp.z, hx_yp.w};
        const float hzxp[4] = {hz.y, hz.z, hz.w, hz_r};
        const float hyxp[4] = {hy_c.y, hy_c.z, hy_c.w, hy_r};
        float exv[4] = {ex.x, ex.y, ex.z, ex.w};
        float eyv[4] = {ey.x, ey.y, ey.z, ey.w};
        float ezv[4] = {ez.x, ez.y, ez.z, ez.w};
#pragma unroll
        for (int k = 0; k < 4; ++k) {
            const bool xk_in = k < xlim;
            const float dhz_dy = y_in ? deriv<EXACT>(hzyp[k], hzv[k], L.dx[1], L.inv_dx[1]) : 0.f;
            const float dhx_dy = y_in ? deriv<EXACT>(hxyp[k], hxv[k], L.dx[1], L.inv_dx[1]) : 0.f;
            const float dhy_dz = z_in ? deriv<EXACT>(hyzp[k], hyv[k], L.dx[2], L.inv_dx[2]) : 0.f;
            const float dhx_dz = z_in ? deriv<EXACT>(hxzp[k], hxv[k], L.dx[2], L.inv_dx[2]) : 0.f;
            const float dhz_dx = xk_in ? deriv<EXACT>(hzxp[k], hzv[k], L.dx[0], L.inv_dx[0]) : 0.f;
            const float dhy_dx = xk_in ? deriv<EXACT>(hyxp[k], hyv[k], L.dx[0], L.inv_dx[0]) : 0.f;
            const float cx = dhz_dy - dhy_dz;
            const float cy = dhx_dz - dhz_dx;
            const float cz = dhy_dx - dhx_dy;
            float jx = 0.f, jy = 0.f, jz = 0.f;
            if (cnt) { const float3 v = forcing_lookup(F, s0, cnt, static_cast<long long>(i) + k, 0); jx = v.x; jy = v.y; jz = v.z; }
            float2 c;
            if (L.table_in_smem) c = s_tab[mk[k]];
            else { const float4 t4 = __ldg(L.table + mk[k]); c = make_float2(t4.x, t4.y); }
            exv[k] = leap<EXACT>(c.x, exv[k], c.y, cx, jx);
            eyv[k] = leap<EXACT>(c.x, eyv[k], c.y, cy, jy);
            ezv[k] = leap<EXACT>(c.x, ezv[k], c.y, cz, jz);
        }
        ex = make_float4(exv[0], exv[1], exv[2], exv[3]);
        ey = make_float4(eyv[0], eyv[1], eyv[2], eyv[3]);
        ez = make_float4(ezv[0], ezv[1], ezv[2], ezv[3]);
        if (act) {
            st4(L.ex + i, ex);
            st4(L.ey + i, ey);
            st4(L.ez + i, ez);
            if (has_probe) {
                for (int s = lower_bound_ll(P.cells, P.n, static_cast<long long>(i)); s < P.n; ++s) {
                    const long long c = __ldg(P.cells + s);
                    if (c >= static_cast<long long>(i) + 4) break;
                    const int k = static_cast<int>(c - static_cast<long long>(i));
                    float* o = P.row + 6 * static_cast<size_t>(__ldg(P.slot + s));
                    o[0] = pick(ex, k); o[1] = pick(ey, k); o[2] = pick(ez, k);
                    o[3] = pick(hx_c, k); o[4] = pick(hy_c, k); o[5] = pick(hz, k);
                }
            }
        }
        hx_c = hx_n; hy_c = hy_n;
    }
}

// ------------------------------------------------------------------------------------------------
// Generic kernels: one thread per cell, any lattice shape (nx % 4 != 0, 1×1×N, 100×100×1 ...).
// ------------------------------------------------------------------------------------------------

template <typename IdxT>
__device__ __forceinline__ int mat_at(const void* base, size_t i) { return static_cast<const IdxT*>(base)[i]; }

template <int EXACT, typename IdxT>
__global__ void __launch_bounds__(256)
k_update_h_generic(const LatticeView L, const ForcingView F, const int z_lo, const int z_hi) {
    const int x = blockIdx.x * blockDim.x + threadIdx.x;
    const int y = blockIdx.y * blockDim.y + threadIdx.y;
    if (x >= L.nx || y >= L.ny) return;
    const size_t sy = static_cast<size_t>(L.nx);
    const size_t sz = sy * static_cast<size_t>(L.ny);
    for (int z = z_lo + blockIdx.z; z < z_hi; z += gridDim.z) {
        const size_t i = static_cast<size_t>(z) * sz + static_cast<size_t>(y) * sy + static_cast<size_t>(x);
        const float ex = L.ex[i], ey = L.ey[i], ez = L.ez[i];
        const bool x_in = x > 0, y_in = y > 0, z_in = (z > 0) || (L.lo_open != 0);
        const float dez_dy = y_in ? deriv<EXACT>(ez, L.ez[i - sy], L.dx[1], L.inv_dx[1]) : 0.f;
        const float dex_dy = y_in ? deriv<EXACT>(ex, L.ex[i - sy], L.dx[1], L.inv_dx[1]) : 0.f;
        const float dey_dz = z_in ? deriv<EXACT>(ey, L.ey[i - sz], L.dx[2], L.inv_dx[2]) : 0.f;
        const float dex_dz = z_in ? deriv<EXACT>(ex, L.ex[i - sz], L.dx[2], L.inv_dx[2]) : 0.f;
        const float dez_dx = x_in ? deriv<EXACT>(ez, L.ez[i - 1], L.dx[0], L.inv_dx[0]) : 0.f;
        const float dey_dx = x_in ? deriv<EXACT>(ey, L.ey[i - 1], L.dx[0], L.inv_dx[0]) : 0.f;
        const float cx = dez_dy - dey_dz;
        const float cy = dex_dz - dez_dx;
        const float cz = dey_dx - dex_dy;
        float mx = 0.f, my = 0.f, mz = 0.f;
        if (F.n) {
            const int s = lower_bound_ll(F.cells, F.n, static_cast<long long>(i));
            if (s < F.n && __ldg(F.cells + s) == static_cast<long long>(i)) {
                mx = __ldg(F.vals + 6 * s + 3); my = __ldg(F.vals + 6 * s + 4); mz = __ldg(F.vals + 6 * s + 5);
            }
        }
        const float4 c = __ldg(L.table + mat_at<IdxT>(L.mat, i));
        L.hx[i] = leap<EXACT>(c.z, L.hx[i], c.w, -cx, mx);
        L.hy[i] = leap<EXACT>(c.z, L.hy[i], c.w, -cy, my);
        L.hz[i] = leap<EXACT>(c.z, L.hz[i], c.w, -cz, mz);
    }
}

template <int EXACT, typename IdxT>
__global__ void __launch_bounds__(256)
k_update_e_generic(const LatticeView L, const ForcingView F, const ProbeView P, const int z_lo, const int z_hi) {
    const int x = blockIdx.x * blockDim.x + threadIdx.x;
    const int y = blockIdx.y * blockDim.y + threadIdx.y;
    if (x >= L.nx || y >= L.ny) return;
    const size_t sy = static_cast<size_t>(L.nx);
    const size_t sz = sy * static_cast<size_t>(L.ny);
    for (int z = z_lo + blockIdx.z; z < z_hi; z += gridDim.z) {
        const size_t i = static_cast<size_t>(z) * sz + static_cast<size_t>(y) * sy + static_cast<size_t>(x);
        const float hx = L.hx[i], hy = L.hy[i], hz = L.hz[i];
        const bool x_in = x + 1 < L.nxl, y_in = y + 1 < L.ny, z_in = (z + 1 < L.nz) || (L.hi_open != 0);
        const float dhz_dy = y_in ? deriv<EXACT>(L.hz[i + sy], hz, L.dx[1], L.inv_dx[1]) : 0.f;
        const float dhx_dy = y_in ? deriv<EXACT>(L.hx[i + sy], hx, L.dx[1], L.inv_dx[1]) : 0.f;
        const float dhy_dz = z_in ? deriv<EXACT>(L.hy[i + sz], hy, L.dx[2], L.inv_dx[2]) : 0.f;
        const float dhx_dz = z_in ? deriv<EXACT>(L.hx[i + sz], hx, L.dx[2], L.inv_dx[2]) : 0.f;
        const float dhz_dx = x_in ? deriv<EXACT>(L.hz[i + 1], hz, L.dx[0], L.inv_dx[0]) : 0.f;
        const float dhy_dx = x_in ? deriv<EXACT>(L.hy[i + 1], hy, L.dx[0], L.inv_dx[0]) : 0.f;
        const float cx = dhz_dy - dhy_dz;
        const float cy = dhx_dz - dhz_dx;
        const float cz = dhy_dx - dhx_dy;
        float jx = 0.f, jy = 0.f, jz = 0.f;
        if (F.n) {
            const int s = lower_bound_ll(F.cells, F.n, static_cast<long long>(i));
            if (s < F.n && __ldg(F.cells + s) == static_cast<long long>(i)) {
                jx = __ldg(F.vals + 6 * s); jy = __ldg(F.vals + 6 * s + 1); jz = __ldg(F.vals + 6 * s + 2);
            }
        }
        const float4 c = __ldg(L.table + mat_at<IdxT>(L.mat, i));
        const float ex = leap<EXACT>(c.x, L.ex[i], c.y, cx, jx);
        const float ey = leap<EXACT>(c.x, L.ey[i], c.y, cy, jy);
        const float ez = leap<EXACT>(c.x, L.ez[i], c.y, cz, jz);
        L.ex[i] = ex; L.ey[i] = ey; L.ez[i] = ez;
        if (P.n) {
            for (int s = lower_bound_ll(P.cells, P.n, static_cast<long long>(i)); s < P.n && __ldg(P.cells + s) == static_cast<long long>(i); ++s) {
                float* o = P.row + 6 * static_cast<size_t>(__ldg(P.slot + s));
                o[0] = ex; o[1] = ey; o[2] = ez; o[3] = hx; o[4] = hy; o[5] = hz;
            }
        }
    }
}

// ------------------------------------------------------------------------------------------------
// Field::field readback / FieldMut (lib.rs:97-127): box <-> interleaved xyz staging, x fastest.
// ------------------------------------------------------------------------------------------------
struct BoxView {
    int x0, y0, z0;     // local coordinates of the box origin
    int bx, by, bz;     // box extents
    int nx, ny;         // lattice row / plane sizes
};

__global__ void k_gather_box(const float* __restrict__ fx, const float* __restrict__ fy, const float* __restrict__ fz,
                             const BoxView B, float* __restrict__ out) {
    const size_t n = static_cast<size_t>(B.bx) * B.by * B.bz;
    for (size_t t = blockIdx.x * static_cast<size_t>(blockDim.x) + threadIdx.x; t < n; t += static_cast<size_t>(gridDim.x) * blockDim.x) {
        const int x = static_cast<int>(t % B.bx);
        const size_t r = t / B.bx;
        const int y = static_cast<int>(r % B.by);
        const int z = static_cast<int>(r / B.by);
        const size_t i = (static_cast<size_t>(z + B.z0) * B.ny + (y + B.y0)) * B.nx + (x + B.x0);
        out[3 * t] = fx[i]; out[3 * t + 1] = fy[i]; out[3 * t + 2] = fz[i];
    }
}

__global__ void k_scatter_box(float* __restrict__ fx, float* __restrict__ fy, float* __restrict__ fz, const BoxView B,
                              const float* __restrict__ in) {
    const size_t n = static_cast<size_t>(B.bx) * B.by * B.bz;
    for (size_t t = blockIdx.x * static_cast<size_t>(blockDim.x) + threadIdx.x; t < n; t += static_cast<size_t>(gridDim.x) * blockDim.x) {
        const int x = static_cast<int>(t % B.bx);
        const size_t r = t / B.bx;
        const int y = static_cast<int>(r % B.by);
        const int z = static_cast<int>(r / B.by);
        const size_t i = (static_cast<size_t>(z + B.z0) * B.ny + (y + B.y0)) * B.nx + (x + B.x0);
        fx[i] = in[3 * t]; fy[i] = in[3 * t + 1]; fz[i] = in[3 * t + 2];
    }
}

// Material ids cross the ABI as rows of `nxl` u32 values; on the device a row holds `pitch` ≥ nxl narrow entries (pad cells last).
template <typename IdxT>
__global__ void k_widen_ids(const IdxT* __restrict__ in, uint32_t* __restrict__ out, size_t n, size_t nxl, size_t pitch) {
    for (size_t t = blockIdx.x * static_cast<size_t>(blockDim.x) + threadIdx.x; t < n; t += static_cast<size_t>(gridDim.x) * blockDim.x)
        out[t] = in[(t / nxl) * pitch + t % nxl];
}
template <typename IdxT>
__global__ void k_narrow_ids(const uint32_t* __restrict__ in, IdxT* __restrict__ out, size_t n, size_t nxl, size_t pitch) {
    for (size_t t = blockIdx.x * static_cast<size_t>(blockDim.x) + threadIdx.x; t < n; t += static_cast<size_t>(gridDim.x) * blockDim.x)
        out[(t / nxl) * pitch + t % nxl] = static_cast<IdxT>(in[t]);
}
// the pad cells of `rows` rows → `id` (the all-zero coefficient row)
template <typename IdxT>
__global__ void k_fill_pad_ids(IdxT* __restrict__ out, size_t nxl, size_t pitch, size_t rows, uint32_t id) {
    const size_t w = pitch - nxl, n = rows * w;
    for (size_t t = blockIdx.x * static_cast<size_t>(blockDim.x) + threadIdx.x; t < n; t += static_cast<size_t>(gridDim.x) * blockDim.x)
        out[(t / w) * pitch + nxl + t % w] = static_cast<IdxT>(id);
}

// ------------------------------------------------------------------------------------------------
// Projection (cpu/project.rs:117-165, project.wgsl:46-75): one thread per pixel, RGBA8 out.
// ------------------------------------------------------------------------------------------------
struct ProjectView {
    const float* fx; const float* fy; const float* fz;  // local plane 0 of the sampled field
    int nx, ny, nz_global, z_begin, nz_local;   // nx = row pitch
    int nxl;                                    // lattice cells of a row (what the projection scales by)
    float proj[16];   // column-major
    float cmap[16];   // column-major
    int mode;
    int w, h;
};

__device__ __forceinline__ float srgb_oetf(float c) { return c <= 0.0031308f ? 12.92f * c : 1.055f * powf(c, 1.0f / 2.4f) - 0.055f; }
__device__ __forceinline__ unsigned char unit_u8(float c) { return static_cast<unsigned char>(fminf(255.f, fmaxf(0.f, roundf(c * 255.0f)))); }

__global__ void k_project_rgba8(const ProjectView V, uchar4* __restrict__ out) {
    const int px = blockIdx.x * blockDim.x + threadIdx.x;
    const int py = blockIdx.y * blockDim.y + threadIdx.y;
    if (px >= V.w || py >= V.h) return;
    float u, v;
    if (V.mode == 0) {
        u = static_cast<float>(px) / static_cast<float>(V.w + 1);
        v = 1.0f - static_cast<float>(py) / static_cast<float>(V.h + 1);
    } else {
        u = (static_cast<float>(px) + 0.5f) / static_cast<float>(V.w);
        v = 1.0f - (static_cast<float>(py) + 0.5f) / static_cast<float>(V.h);
    }
    const int dims[3] = {V.nxl, V.ny, V.nz_global};
    int p[3];
#pragma unroll
    for (int r = 0; r < 3; ++r) {
        const float c = __fadd_rn(__fadd_rn(__fadd_rn(__fmul_rn(V.proj[r], u), __fmul_rn(V.proj[4 + r], v)), __fmul_rn(V.proj[8 + r], 0.0f)), V.proj[12 + r]);
        const float n = static_cast<float>(dims[r]);
        float t = roundf(__fmul_rn(c, n - 1.0f));
        if (!(t > 0.0f)) t = 0.0f;
        p[r] = t >= n - 1.0f ? dims[r] - 1 : static_cast<int>(t);
    }
    const int zl = p[2] - V.z_begin;
    if (zl < 0 || zl >= V.nz_local) return;  // another slab owns this pixel
    const size_t i = (static_cast<size_t>(zl) * V.ny + p[1]) * V.nx + p[0];
    const float f[4] = {V.fx[i], V.fy[i], V.fz[i], 1.0f};
    float col[4];
    if (V.mode == 2) {
        col[0] = col[1] = col[2] = col[3] = 0.f;
        const float xz = f[2];
        if (xz > 0.0f) { col[0] = fminf(10.0f * xz, 1.0f); col[3] = fminf(100.0f * xz, 1.0f); }
        else { col[1] = fminf(-10.0f * xz, 1.0f); col[3] = fminf(-100.0f * xz, 1.0f); }
    } else {
#pragma unroll
        for (int r = 0; r < 4; ++r) {
            float acc = 0.f;
#pragma unroll
            for (int c = 0; c < 4; ++c) acc = __fadd_rn(acc, __fmul_rn(V.cmap[4 * c + r], f[c]));
            col[r] = fminf(1.0f, fmaxf(0.0f, acc));
        }
    }
    uchar4 o;
    if (V.mode == 0) {
        o.x = unit_u8(srgb_oetf(col[0])); o.y = unit_u8(srgb_oetf(col[1])); o.z = unit_u8(srgb_oetf(col[2])); o.w = unit_u8(col[3]);
    } else {
        o.x = unit_u8(col[0]); o.y = unit_u8(col[1]); o.z = unit_u8(col[2]); o.w = unit_u8(col[3]);
    }
    out[static_cast<size_t>(py) * V.w + px] = o;
}

}  // namespace fdtd_b200
