// Scene → material-index lattice on the device.  The reference fills its coefficient lattice with one
// host callback per cell, `DomainDescription::material(point)` (cem-solver/src/fdtd/cpu/mod.rs:244-264), which in the
// application is a BVH point query per cell on a single thread: transform the lattice point to world coordinates in f64,
// cast to f32, ask which colliders contain it, take the first hit's material, else the default material
// (cem-app/src/solver/runner.rs:606-631, 909-918).  Here one thread rasterises four consecutive x cells against a short
// list of primitives and writes the material label straight into the u8 / u16 index lattice (or a u32 lattice for the
// host-facing call); 134 M cells × 64 shapes take a few milliseconds instead of minutes.
//
// Every arithmetic step is spelled with round-to-nearest intrinsics in a fixed order (no FMA contraction), so the labels
// are bit-identical to the numpy restatement in oracle/scene_ref.py — material indexing is integer work and must be exact.
#pragma once

#include <cstdint>

namespace fdtd_b200 {

struct SceneShape {          // = fdtd_b200_shape (include/fdtd_b200.h)
    uint32_t kind;           // 0 ball | 1 cuboid | 2 cylinder (axis = local Y, like parry3d's Cylinder)
    uint32_t label;
    float params[3];         // ball: radius,-,- | cuboid: half extents | cylinder: half_height, radius, -
    float rot[9];            // world → local rotation, row-major (the inverse of the entity's rotation)
    float translation[3];    // entity position, world units
    float reserved;
};

struct SceneBox {            // conservative lattice-space bounds of a shape (cells), computed on the host
    int lo[3], hi[3];        // half-open
};

struct SceneView {
    double m[12];            // rows 0..2 of transform_from_solver_to_world, row-major (row 3 is 0 0 0 1)
    const SceneShape* shapes;
    const SceneBox* boxes;
    int n_shapes;
    uint32_t default_label;
    int nx, ny;              // lattice cells per row, rows per plane
    int pitch;               // entries per row of `out` (≥ nx; pad entries are left alone)
    long long z0;            // global z of local plane 0
    long long n_planes;      // planes to rasterise
};

__device__ __forceinline__ bool shape_contains(const SceneShape& s, float px, float py, float pz) {
    const float d0 = __fsub_rn(px, s.translation[0]), d1 = __fsub_rn(py, s.translation[1]), d2 = __fsub_rn(pz, s.translation[2]);
    const float l0 = __fadd_rn(__fadd_rn(__fmul_rn(s.rot[0], d0), __fmul_rn(s.rot[1], d1)), __fmul_rn(s.rot[2], d2));
    const float l1 = __fadd_rn(__fadd_rn(__fmul_rn(s.rot[3], d0), __fmul_rn(s.rot[4], d1)), __fmul_rn(s.rot[5], d2));
    const float l2 = __fadd_rn(__fadd_rn(__fmul_rn(s.rot[6], d0), __fmul_rn(s.rot[7], d1)), __fmul_rn(s.rot[8], d2));
    if (s.kind == 0u) {
        const float r2 = __fadd_rn(__fadd_rn(__fmul_rn(l0, l0), __fmul_rn(l1, l1)), __fmul_rn(l2, l2));
        return r2 <= __fmul_rn(s.params[0], s.params[0]);
    }
    if (s.kind == 1u) return fabsf(l0) <= s.params[0] && fabsf(l1) <= s.params[1] && fabsf(l2) <= s.params[2];
    const float rr = __fadd_rn(__fmul_rn(l0, l0), __fmul_rn(l2, l2));
    return fabsf(l1) <= s.params[0] && rr <= __fmul_rn(s.params[1], s.params[1]);
}

// grid-stride over quads of x cells; OutT = uint8_t / uint16_t (device index) or uint32_t (host-facing labels)
template <typename OutT>
__global__ void __launch_bounds__(256) k_rasterize_scene(const SceneView V, OutT* __restrict__ out) {
    extern __shared__ __align__(16) unsigned char smem_scene[];
    SceneShape* s_shapes = reinterpret_cast<SceneShape*>(smem_scene);
    SceneBox* s_boxes = reinterpret_cast<SceneBox*>(s_shapes + V.n_shapes);
    for (int i = threadIdx.x; i < V.n_shapes; i += blockDim.x) { s_shapes[i] = V.shapes[i]; s_boxes[i] = V.boxes[i]; }
    __syncthreads();

    const int qx = (V.nx + 3) >> 2;
    const long long rows = static_cast<long long>(V.ny) * V.n_planes;
    const long long n_quads = rows * qx;
    for (long long q = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; q < n_quads;
         q += static_cast<long long>(gridDim.x) * blockDim.x) {
        const long long row = q / qx;
        const int x0 = static_cast<int>(q - row * qx) << 2;
        const int zl = static_cast<int>(row / V.ny);
        const int y = static_cast<int>(row - static_cast<long long>(zl) * V.ny);
        const long long zg = V.z0 + zl;
        const int z = static_cast<int>(zg);
        const double fy = static_cast<double>(y), fz = static_cast<double>(zg);
        uint32_t lab[4];
#pragma unroll
        for (int c = 0; c < 4; ++c) {
            const int x = x0 + c;
            const double fx = static_cast<double>(x);
            // (M·(x, y, z, 1)) row by row in f64, then `try_cast::<f32>()` (runner.rs:909-918)
            float p[3];
#pragma unroll
            for (int r = 0; r < 3; ++r) {
                const double w = __dadd_rn(__dadd_rn(__dadd_rn(__dmul_rn(V.m[4 * r], fx), __dmul_rn(V.m[4 * r + 1], fy)),
                                                     __dmul_rn(V.m[4 * r + 2], fz)), V.m[4 * r + 3]);
                p[r] = __double2float_rn(w);
            }
            uint32_t l = V.default_label;
            for (int s = 0; s < V.n_shapes; ++s) {
                const SceneBox& b = s_boxes[s];
                if (x < b.lo[0] || x >= b.hi[0] || y < b.lo[1] || y >= b.hi[1] || z < b.lo[2] || z >= b.hi[2]) continue;
                if (shape_contains(s_shapes[s], p[0], p[1], p[2])) { l = s_shapes[s].label; break; }   // first match wins
            }
            lab[c] = l;
        }
        const long long base = (static_cast<long long>(zl) * V.ny + y) * V.pitch + x0;
        if (((V.nx | V.pitch) & 3) == 0) {   // whole quad inside the row and aligned: one vector store
            if constexpr (sizeof(OutT) == 1) {
                *reinterpret_cast<uint32_t*>(out + base) = lab[0] | (lab[1] << 8) | (lab[2] << 16) | (lab[3] << 24);
                continue;
            } else if constexpr (sizeof(OutT) == 2) {
                *reinterpret_cast<uint2*>(out + base) = make_uint2(lab[0] | (lab[1] << 16), lab[2] | (lab[3] << 16));
                continue;
            } else {
                *reinterpret_cast<uint4*>(out + base) = make_uint4(lab[0], lab[1], lab[2], lab[3]);
                continue;
            }
        }
#pragma unroll
        for (int c = 0; c < 4; ++c)
            if (x0 + c < V.nx) out[base + c] = static_cast<OutT>(lab[c]);
    }
}

}  // namespace fdtd_b200
