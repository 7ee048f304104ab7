// TMA-staged variants of the two half-step kernels (sm_100a): planes are streamed into a shared-memory
// ring by 1-D bulk async copies (`cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes`, the TMA
// engine — SASS UBLKCP) signalled through mbarriers, so a CTA keeps S-1 planes in flight without holding
// them in registers.  Arithmetic, boundary rule, forcing and probes are the helpers of kernels.cuh, so the
// results are bit-identical to the register-streaming kernels.  Semantics: cem-solver/src/fdtd/cpu/mod.rs:
// 416-515, fdtd/wgpu/update.wgsl:56-177.
#pragma once

#include "kernels.cuh"

namespace fdtd_b200 {

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return static_cast<uint32_t>(__cvta_generic_to_shared(p)); }

__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_arrive_expect_tx(uint64_t* bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint64_t* bar) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ bool mbar_try_wait(uint64_t* bar, uint32_t parity) {
    uint32_t ok;
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n"
        "selp.u32 %0, 1, 0, p;\n"
        "}\n"
        : "=r"(ok)
        : "r"(smem_u32(bar)), "r"(parity)
        : "memory");
    return ok != 0;
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
    while (!mbar_try_wait(bar, parity)) {}
}
// global → shared bulk copy of `bytes` (multiple of 16, both addresses 16-byte aligned), completing on `bar`
__device__ __forceinline__ void bulk_g2s(void* dst_smem, const void* src_gmem, uint32_t bytes, uint64_t* bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(smem_u32(dst_smem)),
                 "l"(src_gmem), "r"(bytes), "r"(smem_u32(bar))
                 : "memory");
}

// Shared-memory layout of one stage (one z-plane of the CTA's tile).  Rows of components that need an x halo
// are PH = 4*BX + 4 floats wide (one extra float4), the others PW = 4*BX.
template <int BX, int BY>
struct BulkLayout {
    static constexpr int PW = 4 * BX;
    static constexpr int PH = 4 * BX + 4;
};

// ------------------------------------------------------------------------------------------------
// H half-step.  Stage contents: Ex rows y0-1..y0+BY-1 (PW), Ey rows y0.. (PH, halo at the left),
// Ez rows y0-1.. (PH, left halo), Hx,Hy,Hz rows y0.. (PW), material index rows y0.. (PW entries).
// ------------------------------------------------------------------------------------------------
template <int EXACT, typename IdxT, int BX, int BY, int S>
__global__ void __launch_bounds__(BX * BY)
k_update_h_bulk(const LatticeView L, const ForcingView F, const int z_lo, const int z_hi, const int zc) {
    using LO = BulkLayout<BX, BY>;
    constexpr int PW = LO::PW, PH = LO::PH;
    constexpr int N_EX = (BY + 1) * PW, N_EY = BY * PH, N_EZ = (BY + 1) * PH, N_H = BY * PW;
    constexpr int STAGE_FLOATS = N_EX + N_EY + N_EZ + 3 * N_H;
    constexpr int MAT_BYTES = BY * PW * static_cast<int>(sizeof(IdxT));
    constexpr int STAGE_BYTES = STAGE_FLOATS * 4 + MAT_BYTES;
    static_assert(STAGE_BYTES % 16 == 0, "stage must stay 16-byte aligned");

    extern __shared__ __align__(128) unsigned char smem_raw[];
    uint64_t* full = reinterpret_cast<uint64_t*>(smem_raw);                 // S mbarriers
    unsigned char* stage0 = smem_raw + 128;
    float2* s_tab = reinterpret_cast<float2*>(stage0 + static_cast<size_t>(S) * STAGE_BYTES);

    const int tid = threadIdx.y * BX + threadIdx.x;
    if (tid == 0) {
        for (int s = 0; s < S; ++s) mbar_init(full + s, 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    }
    if (L.table_in_smem) {
        for (int t = tid; t < L.n_table; t += BX * BY) {
            const float4 c = L.table[t];
            s_tab[t] = make_float2(c.z, c.w);
        }
    }
    __syncthreads();

    const int x0 = blockIdx.x * (4 * BX);
    const int y0 = blockIdx.y * BY;
    const int za = z_lo + blockIdx.z * zc;
    const int zb = min(za + zc, z_hi);
    const int n_planes = zb - za;
    bool has_src = false;
    if (F.n) has_src = __syncthreads_or(any_cell_in_box(F.cells, F.n, tid, BX * BY, x0, x0 + 4 * BX, y0, y0 + BY, za, zb, L.nx,
                                                        static_cast<long long>(L.nx) * L.ny)) != 0;
    if (n_planes <= 0) return;
    const int x = x0 + 4 * threadIdx.x;
    const int y = y0 + threadIdx.y;
    const bool act = (x < L.nx) && (y < L.ny);
    const unsigned lane = threadIdx.x & 31u;
    const size_t sy = static_cast<size_t>(L.nx);
    const size_t sz = sy * static_cast<size_t>(L.ny);

    const int cols = min(4 * BX, L.nx - x0);          // valid cells of the tile row
    const int rows = min(BY, L.ny - y0);
    const bool halo_x = x0 > 0;
    const bool halo_y = y0 > 0;
    const uint32_t row_bytes = static_cast<uint32_t>(cols) * 4u;
    const uint32_t stage_tx = row_bytes * static_cast<uint32_t>((rows + (halo_y ? 1 : 0)) * 2 + rows * 4) +
                              (halo_x ? 16u * static_cast<uint32_t>(rows + rows + (halo_y ? 1 : 0)) : 0u) +
                              static_cast<uint32_t>(cols * sizeof(IdxT)) * static_cast<uint32_t>(rows);

    auto issue = [&](int plane, int slot) {
        unsigned char* st = stage0 + static_cast<size_t>(slot) * STAGE_BYTES;
        float* s_ex = reinterpret_cast<float*>(st);
        float* s_ey = s_ex + N_EX;
        float* s_ez = s_ey + N_EY;
        float* s_hx = s_ez + N_EZ;
        float* s_hy = s_hx + N_H;
        float* s_hz = s_hy + N_H;
        unsigned char* s_mat = reinterpret_cast<unsigned char*>(s_hz + N_H);
        uint64_t* bar = full + slot;
        mbar_arrive_expect_tx(bar, stage_tx);
        const size_t base = static_cast<size_t>(plane) * sz + static_cast<size_t>(y0) * sy + static_cast<size_t>(x0);
        const int r_lo = halo_y ? -1 : 0;
        for (int r = r_lo; r < rows; ++r) {            // smem row index r + 1 for the y-halo components
            const long long g = static_cast<long long>(base) + static_cast<long long>(r) * static_cast<long long>(sy);
            bulk_g2s(s_ex + (r + 1) * PW, L.ex + g, row_bytes, bar);
            bulk_g2s(s_ez + (r + 1) * PH + (halo_x ? 0 : 4), L.ez + g - (halo_x ? 4 : 0), row_bytes + (halo_x ? 16u : 0u), bar);
        }
        for (int r = 0; r < rows; ++r) {
            const size_t g = base + static_cast<size_t>(r) * sy;
            bulk_g2s(s_ey + r * PH + (halo_x ? 0 : 4), L.ey + g - (halo_x ? 4 : 0), row_bytes + (halo_x ? 16u : 0u), bar);
            bulk_g2s(s_hx + r * PW, L.hx + g, row_bytes, bar);
            bulk_g2s(s_hy + r * PW, L.hy + g, row_bytes, bar);
            bulk_g2s(s_hz + r * PW, L.hz + g, row_bytes, bar);
            bulk_g2s(s_mat + static_cast<size_t>(r) * PW * sizeof(IdxT), static_cast<const IdxT*>(L.mat) + g,
                     static_cast<uint32_t>(cols * sizeof(IdxT)), bar);
        }
    };

    if (tid == 0) {
        for (int s = 0; s < S && s < n_planes; ++s) issue(za + s, s);
    }

    // the z-1 plane of Ex,Ey is carried in registers; the chunk's first one comes straight from global memory
    size_t i = static_cast<size_t>(za) * sz + static_cast<size_t>(min(y, L.ny - 1)) * sy + static_cast<size_t>(min(x, L.nx - 4));
    const bool y_in = y > 0;
    const bool x_in = x > 0;
    bool z_in = (za > 0) || (L.lo_open != 0);
    float4 ex_zm = make_float4(0.f, 0.f, 0.f, 0.f), ey_zm = ex_zm;
    if (z_in) { ex_zm = ld4(L.ex + i - sz); ey_zm = ld4(L.ey + i - sz); }

    const int ty = threadIdx.y, tx4 = 4 * threadIdx.x;
    for (int it = 0; it < n_planes; ++it, i += sz) {
        const int slot = it % S;
        mbar_wait(full + slot, static_cast<uint32_t>((it / S) & 1));
        const unsigned char* st = stage0 + static_cast<size_t>(slot) * STAGE_BYTES;
        const float* s_ex = reinterpret_cast<const float*>(st);
        const float* s_ey = s_ex + N_EX;
        const float* s_ez = s_ey + N_EY;
        const float* s_hx = s_ez + N_EZ;
        const float* s_hy = s_hx + N_H;
        const float* s_hz = s_hy + N_H;
        const unsigned char* s_mat = reinterpret_cast<const unsigned char*>(s_hz + N_H);

        const float4 ex = ld4(s_ex + (ty + 1) * PW + tx4);
        const float4 ey = ld4(s_ey + ty * PH + 4 + tx4);
        const float4 ez = ld4(s_ez + (ty + 1) * PH + 4 + tx4);
        float4 hx = ld4(s_hx + ty * PW + tx4), hy = ld4(s_hy + ty * PW + tx4), hz = ld4(s_hz + ty * PW + tx4);
        const float4 ex_ym = ld4(s_ex + ty * PW + tx4);
        const float4 ez_ym = ld4(s_ez + ty * PH + 4 + tx4);
        int mk[4];
        Idx4<IdxT>::load(s_mat, static_cast<size_t>(ty) * PW + tx4, mk);
        if (!act) { mk[0] = mk[1] = mk[2] = mk[3] = 0; }  // surplus threads read shared memory no copy has filled
        float ey_l = __shfl_up_sync(0xffffffffu, ey.w, 1);
        float ez_l = __shfl_up_sync(0xffffffffu, ez.w, 1);
        if (lane == 0) { ey_l = s_ey[ty * PH + 3 + tx4]; ez_l = s_ez[(ty + 1) * PH + 3 + tx4]; }
        __syncthreads();                                  // every thread has drained this slot
        if (tid == 0 && it + S < n_planes) issue(za + it + S, slot);

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

// ------------------------------------------------------------------------------------------------
// E half-step.  Stage contents: Ex,Ey,Ez rows y0.. (PW), Hx rows y0..y0+BY (PW), Hy rows y0.. (PH, halo at
// the right), Hz rows y0..y0+BY (PH, right halo), material index.  Hx,Hy of plane z+1 are read from the
// next stage of the ring (the chunk's last plane reads them from global memory).
// ------------------------------------------------------------------------------------------------
template <int EXACT, typename IdxT, int BX, int BY, int S>
__global__ void __launch_bounds__(BX * BY)
k_update_e_bulk(const LatticeView L, const ForcingView F, const ProbeView P, const int z_lo, const int z_hi, const int zc) {
    using LO = BulkLayout<BX, BY>;
    constexpr int PW = LO::PW, PH = LO::PH;
    constexpr int N_E = BY * PW, N_HX = (BY + 1) * PW, N_HY = BY * PH, N_HZ = (BY + 1) * PH;
    constexpr int STAGE_FLOATS = 3 * N_E + N_HX + N_HY + N_HZ;
    constexpr int MAT_BYTES = BY * PW * static_cast<int>(sizeof(IdxT));
    constexpr int STAGE_BYTES = STAGE_FLOATS * 4 + MAT_BYTES;
    static_assert(STAGE_BYTES % 16 == 0, "stage must stay 16-byte aligned");

    extern __shared__ __align__(128) unsigned char smem_raw[];
    uint64_t* full = reinterpret_cast<uint64_t*>(smem_raw);
    unsigned char* stage0 = smem_raw + 128;
    float2* s_tab = reinterpret_cast<float2*>(stage0 + static_cast<size_t>(S) * STAGE_BYTES);

    const int tid = threadIdx.y * BX + threadIdx.x;
    if (tid == 0) {
        for (int s = 0; s < S; ++s) mbar_init(full + s, 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    }
    if (L.table_in_smem) {
        for (int t = tid; t < L.n_table; t += BX * BY) {
            const float4 c = L.table[t];
            s_tab[t] = make_float2(c.x, c.y);
        }
    }
    __syncthreads();

    const int x0 = blockIdx.x * (4 * BX);
    const int y0 = blockIdx.y * BY;
    const int za = z_lo + blockIdx.z * zc;
    const int zb = min(za + zc, z_hi);
    const int n_planes = zb - za;
    bool has_src = false, has_probe = false;
    if (F.n) has_src = __syncthreads_or(any_cell_in_box(F.cells, F.n, tid, BX * BY, x0, x0 + 4 * BX, y0, y0 + BY, za, zb, L.nx,
                                                        static_cast<long long>(L.nx) * L.ny)) != 0;
    if (P.n) has_probe = __syncthreads_or(any_cell_in_box(P.cells, P.n, tid, BX * BY, x0, x0 + 4 * BX, y0, y0 + BY, za, zb, L.nx,
                                                          static_cast<long long>(L.nx) * L.ny)) != 0;
    if (n_planes <= 0) return;
    const int x = x0 + 4 * threadIdx.x;
    const int y = y0 + threadIdx.y;
    const bool act = (x < L.nx) && (y < L.ny);
    const unsigned lane = threadIdx.x & 31u;
    const size_t sy = static_cast<size_t>(L.nx);
    const size_t sz = sy * static_cast<size_t>(L.ny);

    const int cols = min(4 * BX, L.nx - x0);
    const int rows = min(BY, L.ny - y0);
    const bool halo_x = x0 + 4 * BX < L.nx;           // a float4 to the right of the tile exists
    const bool halo_y = y0 + BY < L.ny;               // a row above the tile exists (then rows == BY)
    const uint32_t row_bytes = static_cast<uint32_t>(cols) * 4u;
    const uint32_t stage_tx = row_bytes * static_cast<uint32_t>(6 * rows + (halo_y ? 2 : 0)) +
                              (halo_x ? 16u * static_cast<uint32_t>(2 * rows + (halo_y ? 1 : 0)) : 0u) +
                              static_cast<uint32_t>(cols * sizeof(IdxT)) * static_cast<uint32_t>(rows);

    auto issue = [&](int plane, int slot) {
        unsigned char* st = stage0 + static_cast<size_t>(slot) * STAGE_BYTES;
        float* s_ex = reinterpret_cast<float*>(st);
        float* s_ey = s_ex + N_E;
        float* s_ez = s_ey + N_E;
        float* s_hx = s_ez + N_E;
        float* s_hy = s_hx + N_HX;
        float* s_hz = s_hy + N_HY;
        unsigned char* s_mat = reinterpret_cast<unsigned char*>(s_hz + N_HZ);
        uint64_t* bar = full + slot;
        mbar_arrive_expect_tx(bar, stage_tx);
        const size_t base = static_cast<size_t>(plane) * sz + static_cast<size_t>(y0) * sy + static_cast<size_t>(x0);
        const uint32_t hb = row_bytes + (halo_x ? 16u : 0u);
        for (int r = 0; r < rows; ++r) {
            const size_t g = base + static_cast<size_t>(r) * sy;
            bulk_g2s(s_ex + r * PW, L.ex + g, row_bytes, bar);
            bulk_g2s(s_ey + r * PW, L.ey + g, row_bytes, bar);
            bulk_g2s(s_ez + r * PW, L.ez + g, row_bytes, bar);
            bulk_g2s(s_hy + r * PH, L.hy + g, hb, bar);
            bulk_g2s(s_mat + static_cast<size_t>(r) * PW * sizeof(IdxT), static_cast<const IdxT*>(L.mat) + g,
                     static_cast<uint32_t>(cols * sizeof(IdxT)), bar);
        }
        for (int r = 0; r < rows + (halo_y ? 1 : 0); ++r) {
            const size_t g = base + static_cast<size_t>(r) * sy;
            bulk_g2s(s_hx + r * PW, L.hx + g, row_bytes, bar);
            bulk_g2s(s_hz + r * PH, L.hz + g, hb, bar);
        }
    };

    if (tid == 0) {
        for (int s = 0; s < S && s < n_planes; ++s) issue(za + s, s);
    }

    size_t i = static_cast<size_t>(za) * sz + static_cast<size_t>(min(y, L.ny - 1)) * sy + static_cast<size_t>(min(x, L.nx - 4));
    const bool y_in = y + 1 < L.ny;
    const int xlim = L.nxl - 1 - x;   // cell k of the quad has an x+1 neighbour inside the lattice iff k < xlim (pad cells: kernels.cuh)
    const int ty = threadIdx.y, tx4 = 4 * threadIdx.x;
    float4 hx_c = make_float4(0.f, 0.f, 0.f, 0.f), hy_c = hx_c;

    for (int it = 0; it < n_planes; ++it, i += sz) {
        const int z = za + it;
        const int slot = it % S;
        const unsigned char* st = stage0 + static_cast<size_t>(slot) * STAGE_BYTES;
        const float* s_ex = reinterpret_cast<const float*>(st);
        const float* s_ey = s_ex + N_E;
        const float* s_ez = s_ey + N_E;
        const float* s_hx = s_ez + N_E;
        const float* s_hy = s_hx + N_HX;
        const float* s_hz = s_hy + N_HY;
        const unsigned char* s_mat = reinterpret_cast<const unsigned char*>(s_hz + N_HZ);
        if (it == 0) {
            mbar_wait(full + slot, 0u);
            hx_c = ld4(s_hx + ty * PW + tx4);
            hy_c = ld4(s_hy + ty * PH + tx4);
        }
        const bool z_in = (z + 1 < L.nz) || (L.hi_open != 0);
        float4 hx_n = hx_c, hy_n = hy_c;
        if (it + 1 < n_planes) {
            const int slot_n = (it + 1) % S;
            mbar_wait(full + slot_n, static_cast<uint32_t>(((it + 1) / S) & 1));
            const float* n_ex = reinterpret_cast<const float*>(stage0 + static_cast<size_t>(slot_n) * STAGE_BYTES);
            const float* n_hx = n_ex + 3 * N_E;
            const float* n_hy = n_hx + N_HX;
            hx_n = ld4(n_hx + ty * PW + tx4);
            hy_n = ld4(n_hy + ty * PH + tx4);
        } else if (z_in) {
            hx_n = ld4(L.hx + i + sz);
            hy_n = ld4(L.hy + i + sz);
        }
        const float4 hz = ld4(s_hz + ty * PH + tx4);
        float4 ex = ld4(s_ex + ty * PW + tx4), ey = ld4(s_ey + ty * PW + tx4), ez = ld4(s_ez + ty * PW + tx4);
        const float4 hz_yp = ld4(s_hz + (ty + 1) * PH + tx4);
        const float4 hx_yp = ld4(s_hx + (ty + 1) * PW + tx4);
        int mk[4];
        Idx4<IdxT>::load(s_mat, static_cast<size_t>(ty) * PW + tx4, mk);
        if (!act) { mk[0] = mk[1] = mk[2] = mk[3] = 0; }  // surplus threads read shared memory no copy has filled
        float hz_r = __shfl_down_sync(0xffffffffu, hz.x, 1);
        float hy_r = __shfl_down_sync(0xffffffffu, hy_c.x, 1);
        if (lane == 31) { hz_r = s_hz[ty * PH + tx4 + 4]; hy_r = s_hy[ty * PH + tx4 + 4]; }
        __syncthreads();                                  // slot `it` drained (slot it+1 is still needed)
        if (tid == 0 && it + S < n_planes) issue(za + it + S, slot);

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

// dynamic shared memory of either bulk kernel: mbarriers + S stages + the coefficient table
template <int BX, int BY, int S, typename IdxT>
constexpr size_t bulk_smem_bytes(int n_table_in_smem) {
    return 128 + static_cast<size_t>(S) * (((BY + 1) * (4 * BX) + BY * (4 * BX + 4) + (BY + 1) * (4 * BX + 4) + 3 * BY * (4 * BX)) * 4 +
                                           BY * 4 * BX * sizeof(IdxT)) +
           static_cast<size_t>(n_table_in_smem) * sizeof(float2);
}

}  // namespace fdtd_b200
