// CPML auxiliary-field passes.  The reference carries the coefficient formulas (cem-solver/src/fdtd/pml.rs:34-86)
// and the storage (cpu/mod.rs:242-255, 649-679) but never runs the recursion (`psi = zeros`, cpu/mod.rs:419-425) — parity unpinned.
//
// The PML is separable: per axis a a profile (b_a[i], c_a[i]) over the coordinate along a; a cell takes part in the pass of
// axis a when c_a at its coordinate is non-zero.  Per such cell and half-step two auxiliary values are advanced,
//     G_ab ← b·G_ab + c·∂F_b/∂x_a      G_ac ← b·G_ac + c·∂F_c/∂x_a        ((a, b, c) cyclic; F = E for the H half-step, H' for E)
// and their part of psi = ±curl(G_x, G_y, G_z) is added to the freshly updated field:
//     H half-step:  H_b += d_b·G_ac,  H_c −= d_b·G_ab          E half-step:  E_b −= c_b·G_ac,  E_c += c_b·G_ab
// (the reference's slot `coeff·(∓curl − source + psi)`, applied as a second pass over the PML slabs only, so the streaming
// kernels of the interior stay untouched and ψ is stored for the slabs alone).
#pragma once

#include "kernels.cuh"

namespace fdtd_b200 {

struct PmlAxisView {
    int n;            // PML coordinates along this axis inside the launch (x, y: all of them; z: those of the plane range)
    const int* pos;   // local coordinate of each, ascending
    const float2* bc; // (b, c) of each for this half-step
    float* g0;        // G_ab   layout: axis x [z][y][k], axis y [z][k][x], axis z [k][y][x]
    float* g1;        // G_ac
    int k0;           // first entry of pos/bc/ψ this launch covers (axis z) — 0 otherwise
};

// FIELD 0: H half-step (backward differences of E), 1: E half-step (forward differences of H').
// `src` are the three components differentiated, `dst` the three components corrected (may belong to different buffer sets).
// EXACT: the arithmetic mode of the half-step kernels (kernels.cuh), so that ψ sees the very derivatives the curl used.
template <int FIELD, int AXIS, typename IdxT, int EXACT>
__global__ void __launch_bounds__(256)
k_pml_axis(const LatticeView L, const PmlAxisView P, const float* __restrict__ s0, const float* __restrict__ s1, const float* __restrict__ s2,
           float* d0, float* d1, float* d2, const int z_lo, const int z_hi) {
    const int r0 = blockIdx.x * blockDim.x + threadIdx.x;
    const int r1 = blockIdx.y * blockDim.y + threadIdx.y;
    const int r2 = blockIdx.z;
    int x, y, z, k;
    size_t q;  // ψ index
    if (AXIS == 0) {
        // the slab is only P.n (≈ 2 × thickness) cells wide: threads run over the flattened (y, k) plane so that no lane idles
        // and the ψ accesses of a warp are contiguous
        const int t = (blockIdx.y * gridDim.x + blockIdx.x) * (blockDim.x * blockDim.y) + threadIdx.y * blockDim.x + threadIdx.x;
        if (t >= P.n * L.ny) return;
        k = t % P.n; y = t / P.n; x = __ldg(P.pos + k); z = z_lo + r2;
        q = (static_cast<size_t>(z) * L.ny + y) * P.n + k;
    } else if (AXIS == 1) {
        if (r0 >= L.nx || r1 >= P.n) return;
        k = r1; x = r0; y = __ldg(P.pos + k); z = z_lo + r2;
        q = (static_cast<size_t>(z) * P.n + k) * L.nx + x;
    } else {
        if (r0 >= L.nx || r1 >= L.ny) return;
        k = P.k0 + r2; x = r0; y = r1; z = __ldg(P.pos + k);
        q = (static_cast<size_t>(k) * L.ny + y) * L.nx + x;
    }
    if (z >= z_hi) return;
    const size_t sy = static_cast<size_t>(L.nx), sz = sy * L.ny;
    const size_t i = static_cast<size_t>(z) * sz + static_cast<size_t>(y) * sy + x;
    const size_t stride = AXIS == 0 ? 1 : (AXIS == 1 ? sy : sz);
    const int coord = AXIS == 0 ? x : (AXIS == 1 ? y : z);
    const int extent = AXIS == 0 ? L.nxl : (AXIS == 1 ? L.ny : L.nz);
    constexpr int B = (AXIS + 1) % 3, C = (AXIS + 2) % 3;
    const float* sb = B == 0 ? s0 : (B == 1 ? s1 : s2);
    const float* sc = C == 0 ? s0 : (C == 1 ? s1 : s2);
    float* db = B == 0 ? d0 : (B == 1 ? d1 : d2);
    float* dc = C == 0 ? d0 : (C == 1 ? d1 : d2);

    // Neumann rule of the update (boundary_condition.rs:29-37): no neighbour → the derivative is zero
    float dfb = 0.f, dfc = 0.f;
    if (FIELD == 0) {
        const bool in = coord > 0 || (AXIS == 2 && L.lo_open);
        if (in) { dfb = deriv<EXACT>(sb[i], sb[i - stride], L.dx[AXIS], L.inv_dx[AXIS]); dfc = deriv<EXACT>(sc[i], sc[i - stride], L.dx[AXIS], L.inv_dx[AXIS]); }
    } else {
        const bool in = coord + 1 < extent || (AXIS == 2 && L.hi_open);
        if (in) { dfb = deriv<EXACT>(sb[i + stride], sb[i], L.dx[AXIS], L.inv_dx[AXIS]); dfc = deriv<EXACT>(sc[i + stride], sc[i], L.dx[AXIS], L.inv_dx[AXIS]); }
    }
    const float2 bc = __ldg(P.bc + k);
    const float g0 = fmaf(bc.x, P.g0[q], bc.y * dfb);
    const float g1 = fmaf(bc.x, P.g1[q], bc.y * dfc);
    P.g0[q] = g0;
    P.g1[q] = g1;
    const float4 co = __ldg(L.table + static_cast<const IdxT*>(L.mat)[i]);
    if (FIELD == 0) {
        db[i] = fmaf(co.w, g1, db[i]);
        dc[i] = fmaf(-co.w, g0, dc[i]);
    } else {
        db[i] = fmaf(-co.y, g1, db[i]);
        dc[i] = fmaf(co.y, g0, dc[i]);
    }
}

// E half-step of the cells of a box, out of place: E_out = c_a·E_in + c_b·(∇×H' − J) with the arithmetic of k_update_e_generic.
// The fused step computes E' from an H' that lacks the ψ correction; after the H pass of the PML this kernel recomputes E' in
// the slabs whose H' changed (reading the untouched old E of the current buffer set, writing the pending set).
struct CellBox { int x0, x1, y0, y1, z0, z1; };

template <int EXACT, typename IdxT>
__global__ void __launch_bounds__(256)
k_update_e_box(const LatticeView L, const ForcingView F, float* __restrict__ ox, float* __restrict__ oy, float* __restrict__ oz, const CellBox B) {
    // threads run over the flattened (y, x) rectangle of the box: narrow x ranges (the slabs on the x faces) leave no lane idle
    const int bw = B.x1 - B.x0;
    const int t = (blockIdx.y * gridDim.x + blockIdx.x) * (blockDim.x * blockDim.y) + threadIdx.y * blockDim.x + threadIdx.x;
    if (t >= bw * (B.y1 - B.y0)) return;
    const int x = B.x0 + t % bw;
    const int y = B.y0 + t / bw;
    const size_t sy = static_cast<size_t>(L.nx);
    const size_t sz = sy * static_cast<size_t>(L.ny);
    for (int z = B.z0 + blockIdx.z; z < B.z1; z += gridDim.z) {
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
        const float4 c = __ldg(L.table + static_cast<const IdxT*>(L.mat)[i]);
        ox[i] = leap<EXACT>(c.x, L.ex[i], c.y, cx, jx);
        oy[i] = leap<EXACT>(c.x, L.ey[i], c.y, cy, jy);
        oz[i] = leap<EXACT>(c.x, L.ez[i], c.y, cz, jz);
    }
}

// E and H of the registered probe cells of planes [z_lo, z_hi) → ring row.  With a PML the half-step kernels' own probe
// writes precede the ψ correction, so the row is rewritten here after it.
__global__ void k_sample_probes(const LatticeView L, const ProbeView P, const int z_lo, const int z_hi) {
    const int t = blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= P.n) return;
    const long long c = __ldg(P.cells + t);
    const long long sz = static_cast<long long>(L.nx) * L.ny;
    const int z = static_cast<int>(c / sz);
    if (z < z_lo || z >= z_hi) return;
    float* o = P.row + 6 * static_cast<size_t>(__ldg(P.slot + t));
    o[0] = L.ex[c]; o[1] = L.ey[c]; o[2] = L.ez[c];
    o[3] = L.hx[c]; o[4] = L.hy[c]; o[5] = L.hz[c];
}

}  // namespace fdtd_b200
