// C ABI of fdtd_b200 (include/fdtd_b200.h): host-side state and kernel launches.
// Reference behaviour being replaced: cem-solver/src/fdtd/cpu/mod.rs (instance, state, update pass,
// field views), cem-solver/src/fdtd/wgpu/mod.rs (fp32 data layout), cem-solver/src/fdtd/util.rs
// (coefficients), cem-solver/src/fdtd/cpu/project.rs + wgpu/project.wgsl (projection).
#include "../../include/fdtd_b200.h"

#include <algorithm>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <map>
#include <mutex>
#include <string>
#include <unordered_map>
#include <vector>

#include "kernels.cuh"
#include "kernels_bulk.cuh"
#include "kernels_fused.cuh"
#include "kernels_scene.cuh"
#include "kernels_pml.cuh"
#include "kernels_resident.cuh"
#include "kernels_2d.cuh"

using namespace fdtd_b200;

namespace {

thread_local std::string g_error;

struct Tune {
    int bx = 0, by = 0, zc = 0;  // 0 = choose automatically
    int variant = -1;            // -1 auto | 0 register-streaming kernels | 1 TMA bulk-copy staged kernels
    int stages = 0;              // ring depth of the staged kernels (0 = default)
    int table_smem = 0;          // fused step: largest coefficient table staged in shared memory (0 = default)
    int l2_promotion = 0;        // fused step: L2 promotion of the tensor maps (0 = default 256 B, 1 none, 2 64 B, 3 128 B)
};

}  // namespace

struct fdtd_b200_solver {
    fdtd_b200_config cfg{};
    uint64_t nx = 0, ny = 0, nz = 0;   // global; nx is the ROW PITCH of every array (fields, index, y/z-slab ψ, halo planes)
    // Row padding: lattices whose nx is not a multiple of 16 would miss the float4 / TMA kernels (16-byte rows of every stream,
    // the index bytes included), and the reference derives nx from ceil(size/dx) (fdtd/mod.rs:24-29), so most user lattices are
    // unaligned.  Such handles store rows of nx = nxl rounded up cells; the nx - nxl pad cells of a row index an extra all-zero
    // coefficient row (table entry pad_id), so they stay exactly 0, and the x+1 Neumann test of the E half-step is made against
    // nxl.  Every cell index that crosses the ABI (sources, probes, boxes, material ids, projection, PML profiles) is in terms
    // of the caller's nxl-wide lattice and converted here.
    uint64_t nxl = 0;                  // the caller's nx
    uint32_t pad_id = 0;               // table row of the pad cells (only when nx != nxl)
    uint32_t n_table_user = 0;         // table rows the caller supplied
    uint64_t z_begin = 0, z_end = 0;   // slab
    uint64_t nzl = 0;                  // owned planes
    uint64_t plane = 0;                // nx*ny
    uint64_t n_local = 0;              // plane*nzl
    int device = 0;
    uint32_t flags = 0;
    bool vec4 = false;
    cudaStream_t own_stream = nullptr, stream = nullptr;

    float* alloc[6] = {nullptr, nullptr, nullptr, nullptr, nullptr, nullptr};  // ghost + planes + ghost
    float* f[6] = {nullptr, nullptr, nullptr, nullptr, nullptr, nullptr};      // plane 0
    float* alt_alloc[6] = {nullptr, nullptr, nullptr, nullptr, nullptr, nullptr};  // ping-pong set of the fused step
    float* g[6] = {nullptr, nullptr, nullptr, nullptr, nullptr, nullptr};
    bool fused = false;
    std::string fused_why_not;      // why this handle is not on the fused step (empty when it is); fdtd_b200_step_path
    FusedMaps fused_maps[2][2];     // TMA tensor maps [tile width: 128 | 64 cells][buffer set] (built at the first fused launch of that width)
    bool fused_maps_ready[2] = {false, false};
    int narrow_tiles = -1;          // plain fused launches use the 64-cell tile (decided once per handle: use_narrow_tiles)
    int resident = -1;              // -1 not decided | 0 no | 1 one thread-block cluster | 2 cooperative grid (k_steps_resident)
    std::string resident_why_not;
    unsigned long long* d_sync = nullptr;      // grid-barrier slots of the resident kernel (one per CTA), epochs keep counting across launches
    unsigned long long sync_epoch = 0;
    int resident_ctas = 0, resident_cpt = 1;   // CTAs of the resident launch, cells per thread
    int step2d = -1;                // -1 not decided | 0 no | 1 the y-marching fused kernel for nz == 1 (kernels_2d.cuh)
    int step2d_slots = 0;           // CTAs of k_step_2d resident on the device at once (occupancy × SMs)
    int zc_override = 0;            // planes per CTA march forced for the next fused launches (fdtd_b200_p2p_step's chunk-aligned split)
    int n_sm = 0;                   // SMs of the device (queried at the first fused launch)
    int sm_smem_bytes = 0, sm_smem_reserved = 0;   // shared memory per SM / reserved per CTA (queried at the first fused launch)
    int cur_set = 0;                // which buffer set f[] currently points at
    void* mat = nullptr;
    int idx_bytes = 1;
    float4* table = nullptr;
    uint32_t n_table = 0;

    // forcing table: rows consumed one per step
    long long* d_src_cells = nullptr;  // sorted local indices
    float* d_src_vals = nullptr;       // [rows][n_src][6]
    uint32_t n_src = 0;
    uint64_t src_rows = 0, src_row = 0;
    size_t src_cells_cap = 0, src_vals_cap = 0;
    // per-step protocol (one update pass = set_sources → step(1) → read_probes, runner.rs:522-526) without host round trips:
    // the step's forcing row goes through a ring of pinned slots by cudaMemcpyAsync in stream order (no synchronize while
    // the cell set is unchanged), and a one-step fdtd_b200_step mirrors its probe row into pinned host memory
    static constexpr int PIN_SLOTS = 8;
    std::vector<long long> src_cells_host;   // what d_src_cells holds
    std::vector<uint64_t> src_cells_raw;     // the caller's cell list that produced it (per-step fast path: compared, not rebuilt)
    std::vector<uint32_t> src_perm;          // column of the caller's list behind each device entry
    float* pin_src = nullptr;                // [PIN_SLOTS][pin_src_floats]
    size_t pin_src_floats = 0;
    cudaEvent_t pin_src_ev[PIN_SLOTS] = {};
    int pin_src_next = 0;
    float* pin_probe = nullptr;              // [PIN_SLOTS][n_probes_total * 6]
    uint64_t pin_probe_tick[PIN_SLOTS] = {}; // tick + 1 of the row a slot mirrors (0: none)
    cudaEvent_t pin_probe_ev[PIN_SLOTS] = {}; // recorded behind the slot's copy: reading a mirrored row waits for this event only,
                                              // so the caller can read step s while step s + 1 is already running
    size_t pin_probe_floats = 0;

    // probes
    long long* d_probe_cells = nullptr;
    std::vector<long long> probe_cells_host;   // what d_probe_cells holds (sorted local indices)
    int* d_probe_slot = nullptr;
    float* d_probe_ring = nullptr;  // [ring_steps][n_probes_total][6]
    // One-step calls (fdtd_b200_step(h, 1), fdtd_b200_p2p_step(h, 1): the reference's per-step protocol) write their probe row
    // into a second ring in mapped pinned HOST memory (rings of at most 64 MB; hd_probe_ring = its device alias): the kernels
    // store the few values straight over PCIe, a read is an event wait + a host memcpy, and no device→host copy sits between
    // two steps.  Batched steps keep writing the HBM ring (the posted PCIe stores cost ≈ 3 µs per step, measured).
    // row_in_host[r] says which ring holds row r.  FDTD_B200_STAGED_IO=1 disables the host ring.
    float* h_probe_ring = nullptr;
    float* hd_probe_ring = nullptr;
    bool probe_rows_to_host = false;   // set for the duration of a one-step call
    std::vector<unsigned char> row_in_host;
    // one-step forcing: the kernels read the step's values straight from the pinned slot (device alias), no host→device copy
    const float* src_vals_mapped = nullptr;
    int pin_src_inflight = -1;      // slot whose event end_step records behind the step that consumes it
    uint32_t n_probes_total = 0, n_probes_local = 0;
    uint32_t ring_steps = 0;
    uint64_t probe_first_tick = 0;  // tick at registration

    // scratch for read/write_field
    float* d_stage = nullptr;
    size_t stage_floats = 0;

    uint64_t tick = 0;
    double time = 0.0;
    uint64_t launches = 0;
    // per-kernel CUDA-event profiling (bench.py roofline) and a stream timer
    bool profiling = false;
    std::vector<std::pair<cudaEvent_t, cudaEvent_t>> prof_events[2];  // [0]=H launches, [1]=E launches
    uint64_t prof_cells[2] = {0, 0};
    cudaEvent_t timer_start = nullptr, timer_stop = nullptr;
    Tune tune;
    std::string err;

    // CPML (fdtd_b200_set_pml): per axis the owned PML coordinates, their (b, c) and the ψ arrays of the slabs
    struct PmlAxis {
        int n = 0;                     // PML coordinates along the axis owned by this handle
        std::vector<int> pos;          // local coordinates, ascending (host copy of d_pos)
        std::vector<char> h_active;    // c of the H half-step is non-zero at pos[k]
        int* d_pos = nullptr;
        float2* d_bc[2] = {nullptr, nullptr};  // [0] H half-step, [1] E half-step
        float* psi[2][2] = {{nullptr, nullptr}, {nullptr, nullptr}};  // [half-step][G_ab, G_ac]
        float* psi_alt[2] = {nullptr, nullptr};  // fused step with the CPML inside the kernel: ψ_H of the pending set
        int* d_map = nullptr;          // … and coordinate → k (index into pos), -1 outside the slabs
        bool ghost = false;            // axis z of a slab: the neighbour's plane above (local coordinate nzl) is a PML coordinate;
                                       // its (b, c) are entry n of d_bc, its map entry is n
        size_t floats = 0;             // length of each ψ array
    } pml[3];
    // fused PML on z-slabs: ψ_H of the upper neighbour's first plane, pushed here by that neighbour (peer-memory transport);
    // two slots by tick parity, each [axis x: G_ab, G_ac | axis y | axis z], plane sizes in ghost_plane[]
    float* psi_ghost = nullptr;
    size_t ghost_plane[3] = {0, 0, 0};
#ifdef FDTD_B200_CTA_TRACE
    unsigned long long* trace_buf = nullptr;
    int trace_ctas = 0, trace_grid[4] = {0, 0, 0, 0};
#endif
    bool has_pml = false;
    int heavy_key[3] = {-1, -1, -1};   // (z_lo, z_hi, zc) the cached dispatch order below was computed for; reset by pml_refresh_mode
    int heavy_val[2][2] = {{0, 0}, {0, 0}};   // FusedPmlView::heavy of the y and z block indices
    bool pml_in_kernel = false;        // the fused kernel's PML instantiation runs the recursion (else: passes over the slabs)

    // NVLink peer-memory halo transport (fdtd_b200_p2p_*)
    struct PeerLink {
        bool present = false, same_process = false;
        uint64_t nzl = 0;
        float* field[2][6] = {};            // the neighbour's two buffer sets, each starting at its lower ghost plane
        unsigned char* mat = nullptr;
        unsigned long long* flags = nullptr;  // the neighbour's flag words
        std::vector<void*> opened;          // cudaIpcOpenMemHandle results to close
        float* psi_ghost = nullptr;         // the neighbour's ψ_H receive buffer (lower neighbour only)
    } lower, upper;
    unsigned long long* d_flags = nullptr;  // [0] pushes received from below, [1] from above
    // Sticky timeout marker of the flag waits, in mapped pinned HOST memory (hd_timeout = its device alias): a wait that gives
    // up stores the push count it was waiting for; from then on every wait returns at once and every entry point that steps,
    // reads or writes state fails with FDTD_B200_ECUDA (p2p_poisoned) — a lagging or dead neighbour never yields fields or
    // probe rows computed on stale ghost planes.
    volatile unsigned long long* h_timeout = nullptr;
    unsigned long long* hd_timeout = nullptr;
    bool halos_dirty = false;               // reset / set_time / write_field / write_pml_state since the last refresh: the neighbours'
                                            // ghost planes (and psi receive slots) are stale, fdtd_b200_p2p_step refuses to run
    uint64_t halo_seq = 0;                  // pushes made to (and expected from) each neighbour so far
    bool p2p = false, p2p_mat_sent = false;
    cudaStream_t bstream = nullptr;         // high-priority stream of the boundary planes and the pushes
    cudaEvent_t ev_main = nullptr, ev_bnd = nullptr, ev_push = nullptr;
};

namespace {

int fail(fdtd_b200_handle h, int code, const std::string& msg) {
    if (h) h->err = msg;
    g_error = msg;
    return code;
}

#define CU_TRY(h, expr)                                                                                    \
    do {                                                                                                   \
        cudaError_t e__ = (expr);                                                                          \
        if (e__ != cudaSuccess) {                                                                          \
            (void)cudaGetLastError();                                                                      \
            return fail((h), e__ == cudaErrorMemoryAllocation ? FDTD_B200_ENOMEM : FDTD_B200_ECUDA,        \
                        std::string(#expr) + ": " + cudaGetErrorString(e__));                              \
        }                                                                                                  \
    } while (0)

struct MatKey {
    uint64_t w[4];
    bool operator==(const MatKey& o) const { return std::memcmp(w, o.w, sizeof w) == 0; }
};
struct MatKeyHash {
    size_t operator()(const MatKey& k) const {
        uint64_t h = 0x9E3779B97F4A7C15ull;
        for (uint64_t v : k.w) { h ^= v + 0x9E3779B97F4A7C15ull + (h << 6) + (h >> 2); }
        return static_cast<size_t>(h);
    }
};
MatKey key_of(const fdtd_b200_material& m) {
    MatKey k;
    static_assert(sizeof(fdtd_b200_material) == sizeof(k.w), "material is four doubles");
    std::memcpy(k.w, &m, sizeof k.w);
    return k;
}

bool c_or_d(double perm, double sigma, double dt, double* a, double* b) {
    // fdtd/util.rs:104-116
    const double x = 0.5 * sigma * dt / perm;
    *a = (1.0 - x) / (1.0 + x);
    *b = dt / (perm * (1.0 + x));
    return !(std::isnan(*a) || std::isnan(*b));
}

Tune parse_tune() {
    Tune t;
    if (const char* s = std::getenv("FDTD_B200_TUNE")) {
        int a = 0, b = 0, c = 0, d = -1, e = 0, f = 0, g = 0;
        if (std::sscanf(s, "%d,%d,%d,%d,%d,%d,%d", &a, &b, &c, &d, &e, &f, &g) >= 1) { t.bx = a; t.by = b; t.zc = c; t.variant = d; t.stages = e; t.table_smem = f; t.l2_promotion = g; }
    }
    return t;
}

// cudaFuncSetAttribute(MaxDynamicSharedMemorySize) once per kernel and device, not once per launch (≈ 1 µs of host time, which
// shows on lattices whose step is a few µs).  The limit is a property of the function in the device's context, shared by every
// handle of the process, so it is only ever raised.
template <typename K>
int raise_smem_limit(fdtd_b200_handle h, K kernel, size_t smem) {
    static std::mutex mu;
    static std::map<std::pair<int, const void*>, size_t> limit;
    const std::pair<int, const void*> key(h->device, reinterpret_cast<const void*>(kernel));
    std::lock_guard<std::mutex> lock(mu);
    auto it = limit.find(key);
    if (it != limit.end() && it->second >= smem) return FDTD_B200_OK;
    const cudaError_t e = cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(smem));
    if (e != cudaSuccess) { (void)cudaGetLastError(); return -1; }
    limit[key] = smem;
    return FDTD_B200_OK;
}

int set_device(fdtd_b200_handle h) {
    CU_TRY(h, cudaSetDevice(h->device));
    return FDTD_B200_OK;
}

LatticeView lattice_view(fdtd_b200_handle h) {
    LatticeView L{};
    L.ex = h->f[0]; L.ey = h->f[1]; L.ez = h->f[2];
    L.hx = h->f[3]; L.hy = h->f[4]; L.hz = h->f[5];
    L.mat = h->mat;
    L.table = h->table;
    L.n_table = static_cast<int>(h->n_table);
    L.table_in_smem = h->n_table <= 4096 ? 1 : 0;
    L.nx = static_cast<int>(h->nx); L.ny = static_cast<int>(h->ny); L.nz = static_cast<int>(h->nzl);
    L.nxl = static_cast<int>(h->nxl);
    L.lo_open = h->z_begin > 0 ? 1 : 0;
    L.hi_open = h->z_end < h->nz ? 1 : 0;
    for (int a = 0; a < 3; ++a) {
        L.dx[a] = static_cast<float>(h->cfg.dx[a]);  // resolution cast to f32 as wgpu/mod.rs:606-612
        L.inv_dx[a] = 1.0f / L.dx[a];
    }
    return L;
}

ForcingView forcing_view(fdtd_b200_handle h) {
    ForcingView F{};
    if (h->n_src && h->src_row < h->src_rows) {
        F.n = static_cast<int>(h->n_src);
        F.cells = h->d_src_cells;
        F.vals = h->src_vals_mapped ? h->src_vals_mapped : h->d_src_vals + h->src_row * static_cast<uint64_t>(h->n_src) * 6;
    }
    return F;
}

ProbeView probe_view(fdtd_b200_handle h) {
    ProbeView P{};
    if (h->n_probes_local) {
        P.n = static_cast<int>(h->n_probes_local);
        P.cells = h->d_probe_cells;
        P.slot = h->d_probe_slot;
        P.row = ((h->probe_rows_to_host && h->hd_probe_ring) ? h->hd_probe_ring : h->d_probe_ring) +
                (h->tick % h->ring_steps) * static_cast<uint64_t>(h->n_probes_total) * 6;
    }
    return P;
}

// ---- kernel dispatch ---------------------------------------------------------------------------

using HKernel = void (*)(LatticeView, ForcingView, int, int, int);
using EKernel = void (*)(LatticeView, ForcingView, ProbeView, int, int, int);

struct Shape { int bx, by; };
constexpr Shape kShapes[] = {{32, 4}, {64, 2}, {128, 1}, {128, 2}, {128, 4}};
constexpr int kNumShapes = sizeof(kShapes) / sizeof(kShapes[0]);

template <int EXACT, typename IdxT>
HKernel h_kernel(int shape) {
    switch (shape) {
        case 0: return k_update_h_vec4<EXACT, IdxT, 32, 4>;
        case 1: return k_update_h_vec4<EXACT, IdxT, 64, 2>;
        case 2: return k_update_h_vec4<EXACT, IdxT, 128, 1>;
        case 3: return k_update_h_vec4<EXACT, IdxT, 128, 2>;
        default: return k_update_h_vec4<EXACT, IdxT, 128, 4>;
    }
}
template <int EXACT, typename IdxT>
EKernel e_kernel(int shape) {
    switch (shape) {
        case 0: return k_update_e_vec4<EXACT, IdxT, 32, 4>;
        case 1: return k_update_e_vec4<EXACT, IdxT, 64, 2>;
        case 2: return k_update_e_vec4<EXACT, IdxT, 128, 1>;
        case 3: return k_update_e_vec4<EXACT, IdxT, 128, 2>;
        default: return k_update_e_vec4<EXACT, IdxT, 128, 4>;
    }
}

int pick_shape(fdtd_b200_handle h) {
    // 128-thread CTAs were the fastest family in the 512^3 launch-shape sweep (profiles/r01_sweep.md)
    const int quads = static_cast<int>(h->nx / 4);
    int bx = h->tune.bx, by = h->tune.by;
    if (bx == 0) bx = quads >= 128 ? 128 : (quads > 32 ? 64 : 32);
    if (by == 0) by = 128 / bx;
    for (int s = 0; s < kNumShapes; ++s)
        if (kShapes[s].bx == bx && kShapes[s].by == by) return s;
    return bx >= 128 ? 2 : (bx == 64 ? 1 : 0);
}

int pick_zc(fdtd_b200_handle h, int planes) {
    if (h->tune.zc > 0) return h->tune.zc;
    // enough CTAs for several waves over the device's SMs (148 on a B200), but long enough marches that the extra
    // z∓1 prologue plane (2 of 9 streams) stays a small fraction of the traffic
    const Shape sh = kShapes[pick_shape(h)];
    const long long row_blocks = ((static_cast<long long>(h->nx) / 4 + sh.bx - 1) / sh.bx) * ((static_cast<long long>(h->ny) + sh.by - 1) / sh.by);
    int zc = 16;
    while (zc > 4 && row_blocks * ((planes + zc - 1) / zc) < static_cast<long long>(h->n_sm) * 16) zc >>= 1;
    return zc;
}

template <int EXACT, typename IdxT>
int launch_h_t(fdtd_b200_handle h, int z_lo, int z_hi) {
    const LatticeView L = lattice_view(h);
    const ForcingView F = forcing_view(h);
    if (h->vec4) {
        const int s = pick_shape(h);
        const Shape sh = kShapes[s];
        const int zc = pick_zc(h, z_hi - z_lo);
        dim3 block(sh.bx, sh.by, 1);
        dim3 grid((L.nx / 4 + sh.bx - 1) / sh.bx, (L.ny + sh.by - 1) / sh.by, (z_hi - z_lo + zc - 1) / zc);
        const size_t smem = L.table_in_smem ? sizeof(float2) * h->n_table : 0;
        h_kernel<EXACT, IdxT>(s)<<<grid, block, smem, h->stream>>>(L, F, z_lo, z_hi, zc);
    } else {
        dim3 block(64, 4, 1);
        dim3 grid((L.nx + 63) / 64, (L.ny + 3) / 4, std::min(z_hi - z_lo, 65535));
        k_update_h_generic<EXACT, IdxT><<<grid, block, 0, h->stream>>>(L, F, z_lo, z_hi);
    }
    h->launches += 1;
    CU_TRY(h, cudaGetLastError());
    return FDTD_B200_OK;
}

template <int EXACT, typename IdxT>
int launch_e_t(fdtd_b200_handle h, int z_lo, int z_hi) {
    const LatticeView L = lattice_view(h);
    const ForcingView F = forcing_view(h);
    const ProbeView P = probe_view(h);
    if (h->vec4) {
        const int s = pick_shape(h);
        const Shape sh = kShapes[s];
        const int zc = pick_zc(h, z_hi - z_lo);
        dim3 block(sh.bx, sh.by, 1);
        dim3 grid((L.nx / 4 + sh.bx - 1) / sh.bx, (L.ny + sh.by - 1) / sh.by, (z_hi - z_lo + zc - 1) / zc);
        const size_t smem = L.table_in_smem ? sizeof(float2) * h->n_table : 0;
        e_kernel<EXACT, IdxT>(s)<<<grid, block, smem, h->stream>>>(L, F, P, z_lo, z_hi, zc);
    } else {
        dim3 block(64, 4, 1);
        dim3 grid((L.nx + 63) / 64, (L.ny + 3) / 4, std::min(z_hi - z_lo, 65535));
        k_update_e_generic<EXACT, IdxT><<<grid, block, 0, h->stream>>>(L, F, P, z_lo, z_hi);
    }
    h->launches += 1;
    CU_TRY(h, cudaGetLastError());
    return FDTD_B200_OK;
}


// ---- TMA bulk-copy staged kernels (kernels_bulk.cuh) ------------------------------------------------
struct BulkShape { int bx, by, s; };
constexpr BulkShape kBulkShapes[] = {{128, 1, 4}, {128, 2, 3}, {128, 2, 4}, {128, 4, 2}, {128, 4, 3}, {32, 4, 4}};
constexpr int kNumBulkShapes = sizeof(kBulkShapes) / sizeof(kBulkShapes[0]);

bool bulk_eligible(fdtd_b200_handle h) {
    // bulk copies need 16-byte aligned rows of every stream, including the 1- or 2-byte material index
    return h->vec4 && (h->nx * h->idx_bytes) % 16 == 0;
}

int pick_bulk_shape(fdtd_b200_handle h) {
    int bx = h->tune.bx, by = h->tune.by, st = h->tune.stages;
    if (bx == 0) bx = h->nx >= 512 ? 128 : 32;
    if (bx != 128) { bx = 32; }
    if (by == 0) by = bx == 128 ? 2 : 4;
    if (st == 0) st = bx == 128 ? (by == 1 ? 4 : 3) : 4;
    for (int s = 0; s < kNumBulkShapes; ++s)
        if (kBulkShapes[s].bx == bx && kBulkShapes[s].by == by && kBulkShapes[s].s == st) return s;
    return bx == 128 ? 1 : 5;
}

template <int EXACT, typename IdxT, int BX, int BY, int S>
int launch_bulk(fdtd_b200_handle h, bool e_step, int z_lo, int z_hi) {
    const LatticeView L = lattice_view(h);
    const ForcingView F = forcing_view(h);
    int zc = h->tune.zc;
    if (zc <= 0) {
        // long marches amortise the ring's fill latency; keep >= ~3 waves of CTAs over the 148 SMs
        const long long tiles = static_cast<long long>((L.nx + 4 * BX - 1) / (4 * BX)) * ((L.ny + BY - 1) / BY);
        zc = 64;
        while (zc > 8 && tiles * ((z_hi - z_lo + zc - 1) / zc) < static_cast<long long>(h->n_sm) * 2 * 3) zc >>= 1;
    }
    dim3 block(BX, BY, 1);
    dim3 grid((L.nx + 4 * BX - 1) / (4 * BX), (L.ny + BY - 1) / BY, (z_hi - z_lo + zc - 1) / zc);
    const size_t smem = bulk_smem_bytes<BX, BY, S, IdxT>(L.table_in_smem ? L.n_table : 0);
    if (e_step) {
        const ProbeView P = probe_view(h);
        auto k = k_update_e_bulk<EXACT, IdxT, BX, BY, S>;
        if (raise_smem_limit(h, k, smem)) return fail(h, FDTD_B200_ECUDA, "cudaFuncSetAttribute(MaxDynamicSharedMemorySize) failed");
        k<<<grid, block, smem, h->stream>>>(L, F, P, z_lo, z_hi, zc);
    } else {
        auto k = k_update_h_bulk<EXACT, IdxT, BX, BY, S>;
        if (raise_smem_limit(h, k, smem)) return fail(h, FDTD_B200_ECUDA, "cudaFuncSetAttribute(MaxDynamicSharedMemorySize) failed");
        k<<<grid, block, smem, h->stream>>>(L, F, z_lo, z_hi, zc);
    }
    h->launches += 1;
    CU_TRY(h, cudaGetLastError());
    return FDTD_B200_OK;
}

template <int EXACT, typename IdxT>
int launch_bulk_shape(fdtd_b200_handle h, bool e_step, int z_lo, int z_hi) {
    switch (pick_bulk_shape(h)) {
        case 0: return launch_bulk<EXACT, IdxT, 128, 1, 4>(h, e_step, z_lo, z_hi);
        case 1: return launch_bulk<EXACT, IdxT, 128, 2, 3>(h, e_step, z_lo, z_hi);
        case 2: return launch_bulk<EXACT, IdxT, 128, 2, 4>(h, e_step, z_lo, z_hi);
        case 3: return launch_bulk<EXACT, IdxT, 128, 4, 2>(h, e_step, z_lo, z_hi);
        case 4: return launch_bulk<EXACT, IdxT, 128, 4, 3>(h, e_step, z_lo, z_hi);
        default: return launch_bulk<EXACT, IdxT, 32, 4, 4>(h, e_step, z_lo, z_hi);
    }
}

int launch_bulk_any(fdtd_b200_handle h, bool e_step, int z_lo, int z_hi, int mode) {
    if (h->idx_bytes == 1) {
        return mode == 0 ? launch_bulk_shape<0, uint8_t>(h, e_step, z_lo, z_hi)
                         : (mode == 1 ? launch_bulk_shape<1, uint8_t>(h, e_step, z_lo, z_hi) : launch_bulk_shape<2, uint8_t>(h, e_step, z_lo, z_hi));
    }
    return mode == 0 ? launch_bulk_shape<0, uint16_t>(h, e_step, z_lo, z_hi)
                     : (mode == 1 ? launch_bulk_shape<1, uint16_t>(h, e_step, z_lo, z_hi) : launch_bulk_shape<2, uint16_t>(h, e_step, z_lo, z_hi));
}

bool use_bulk(fdtd_b200_handle h) {
    if (!bulk_eligible(h) || h->tune.variant == 0) return false;
    if (h->tune.variant == 1) return true;
    return h->nx >= 128;  // auto: the staged kernels win on wide rows (profiles/r01_sweep_bulk.jsonl)
}


// ---- fused single-pass step (kernels_fused.cuh) --------------------------------------------------------
// Step mode of a handle: the fused single-pass step unless the caller opted out or the lattice is not eligible.  Returns
// an empty string for "fused", else the reason (reported by fdtd_b200_step_path; there are no silent fallbacks).
std::string fused_refusal(uint32_t flags, int tune_variant, uint64_t pitch, int idx_bytes) {
    if (flags & FDTD_B200_FORCE_GENERIC_KERNELS) return "FDTD_B200_FORCE_GENERIC_KERNELS";
    if (flags & FDTD_B200_IN_PLACE_STEP) return "in-place step requested (FDTD_B200_IN_PLACE_STEP)";
    if (tune_variant == 0 || tune_variant == 1) return "FDTD_B200_TUNE selects a two-kernel variant";
    if (pitch % 4 != 0) return "row pitch " + std::to_string(pitch) + " is not a multiple of 4 cells";
    if (pitch < 128) return "row pitch " + std::to_string(pitch) + " < 128 cells (one tile of the fused kernel)";
    if ((pitch * idx_bytes) % 16 != 0) return "index rows are not 16-byte aligned";
    return std::string();
}

// Row pitch of a lattice of nx cells per row (see fdtd_b200_solver::nxl).  Rows of at least 64 cells are padded to the next
// multiple of 16 (≤ 23 % more bytes); rows of 96 … 127 cells to 128, one tile of the fused kernel (≤ 33 % more bytes — a multiple
// of 16 among them too: 112³ ran the two-kernel path at 28.5 µs per step, profiles/r02_exp_narrow_tiles.txt);
// FDTD_B200_NO_ROW_PADDING=1 or the generic-kernel flag keep the caller's nx, and so does a table that has no room for the pad row.
// `cells`: cells of the handle (nx · ny · planes).  Rows of 96 or 112 cells need no padding to be stored, so a lattice of them that
// fits the on-chip multi-step kernels (§3d of DESIGN.md: 148 SMs × 2048 cells) keeps its pitch — padding would push it over that budget.
uint64_t row_pitch(uint64_t nx, uint32_t flags, uint32_t n_table, uint64_t cells) {
    if (nx < 64 || (flags & FDTD_B200_FORCE_GENERIC_KERNELS) || n_table >= 65536) return nx;
    if (const char* e = std::getenv("FDTD_B200_NO_ROW_PADDING")) if (e[0] == '1') return nx;
    if (nx >= 96 && nx < 128) return (nx % 16 == 0 && cells <= 303104) ? nx : 128;
    return (nx + 15) / 16 * 16;
}

using EncodeTiledFn = CUresult (*)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*, const cuuint32_t*,
                                   const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

EncodeTiledFn encode_tiled_fn() {
    static EncodeTiledFn fn = nullptr;
    if (!fn) {
        void* p = nullptr;
        cudaDriverEntryPointQueryResult q;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) == cudaSuccess && q == cudaDriverEntryPointSuccess)
            fn = reinterpret_cast<EncodeTiledFn>(p);
    }
    return fn;
}

// 3-D map (x, y, z) of one lattice array with a box of (bx, by, 1) elements
int make_map(fdtd_b200_handle h, CUtensorMap* out, CUtensorMapDataType dt, size_t elem, void* base, uint32_t bx, uint32_t by, uint64_t extra_planes) {
    EncodeTiledFn enc = encode_tiled_fn();
    if (!enc) return fail(h, FDTD_B200_ECUDA, "cuTensorMapEncodeTiled is not available from this driver");
    const cuuint64_t dims[3] = {h->nx, h->ny, h->nzl + extra_planes};
    const cuuint64_t strides[2] = {h->nx * elem, h->nx * h->ny * elem};
    const cuuint32_t box[3] = {bx, by, 1};
    const cuuint32_t estr[3] = {1, 1, 1};
    // 256-byte promotion measured best (tools/exp_promo.sh, 512³: 1.0388 ms against 1.0450 with 128 B or none, 1.0470 with 64 B;
    // 1024×1024×128: 1.0558 against 1.0670)
    const CUtensorMapL2promotion promo = h->tune.l2_promotion == 1 ? CU_TENSOR_MAP_L2_PROMOTION_NONE
                                         : h->tune.l2_promotion == 2 ? CU_TENSOR_MAP_L2_PROMOTION_L2_64B
                                         : h->tune.l2_promotion == 3 ? CU_TENSOR_MAP_L2_PROMOTION_L2_128B : CU_TENSOR_MAP_L2_PROMOTION_L2_256B;
    const CUresult r = enc(out, dt, 3, base, dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE,
                           promo, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) return fail(h, FDTD_B200_ECUDA, "cuTensorMapEncodeTiled failed with code " + std::to_string(static_cast<int>(r)));
    return FDTD_B200_OK;
}

// Planes per CTA march of the fused kernel.  A march of zc planes costs zc + c plane-times (the extra old-E plane, the ring fill
// and the recomputed H' plane above the chunk), the CTAs run in waves of 2 per SM, and the last wave leaves the GPU partly idle for
// a fraction a of a march: cost(zc) = (ceil(tiles · chunks / slots) + a) · (zc + c) with balanced chunks.  Round 1 fitted a = 0.43,
// c = 2.5 (profiles/r01_exp_zc_midsize.txt) — with the CTAs that hold a source or a probe running up to 29 % longer than the
// others, which is what made a wave's end ragged.  With the per-plane source / probe masks in the kernel the same sweep over six
// lattices fits a = 0.12, c = 2.0 and the model picks the measured best march in each (profiles/r02_exp_cta_trace.txt).
int pick_fused_zc(fdtd_b200_handle h, long long tiles, int planes, int min_chunks = 1, int max_zc = 64) {
    if (h->n_sm == 0) {
        if (cudaDeviceGetAttribute(&h->n_sm, cudaDevAttrMultiProcessorCount, h->device) != cudaSuccess || h->n_sm <= 0) { (void)cudaGetLastError(); h->n_sm = 148; }
    }
    const double slots = 2.0 * h->n_sm;
    // every march length up to the cap (longer marches measured slower: 128 planes −6 %, profiles/r01_exp_plane_shape.txt).  Round 1
    // tried a coarse list (1, 2, 4, 6, 8, 12, 16, 20, …): on small lattices it misses the length that fills one wave — 128³ has 19
    // tiles, 12 planes make 209 CTAs for 296 slots, 9 planes make 285.
    int best = std::max(1, std::min(planes, max_zc));
    double best_cost = 1e300;
    for (int zc = 1; zc <= max_zc; ++zc) {
        if (zc > planes && zc != 1) { if (best_cost < 1e300) break; }
        const int chunks = (planes + zc - 1) / zc;
        if (chunks < min_chunks) continue;                // the peer-memory split wants a chunk per boundary plus an interior
        const int zce = (planes + chunks - 1) / chunks;   // balanced
        const double cost = (std::ceil(static_cast<double>(tiles) * chunks / slots) + 0.12) * (zce + 2.0);   // constants: profiles/r02_exp_cta_trace.txt
        if (cost < best_cost * (1.0 - 1e-9)) { best_cost = cost; best = zce; }
    }
    return best;
}

// Launch of a step kernel with programmatic stream serialization: if the previous operation of the stream is a kernel that
// signals launch_dependents (the step kernels do, kernels.cuh), this launch's CTAs may become resident before that grid has
// drained; they block in griddepcontrol.wait until it has.  FDTD_B200_NO_PDL=1 launches plainly (the A side of the A/B).
// `allow`: false for slabs with a neighbour — their step is a boundary launch (or two) on the priority stream and an interior launch
// on the main stream, ordered by events; letting the second boundary launch overlap the first measured 2.3 % slower at N = 4
// (493 against 505 Gcells/s, same box, profiles/r02_exp_pdl.txt).
template <typename... KArgs, typename... Args>
cudaError_t launch_step_kernel(void (*k)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t st, bool allow, Args&&... args) {
    static const bool enabled = [] { const char* e = std::getenv("FDTD_B200_NO_PDL"); return !(e && e[0] == '1'); }();
    const bool pdl = enabled && allow;
    cudaLaunchConfig_t cfg{};
    cfg.gridDim = grid; cfg.blockDim = block; cfg.dynamicSmemBytes = smem; cfg.stream = st;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr; cfg.numAttrs = pdl ? 1 : 0;
    return cudaLaunchKernelEx(&cfg, k, std::forward<Args>(args)...);
}

// Tile width of the plain fused launches.  The 128-cell tile (one warp per thread row) wastes the lanes of a last tile column
// that is mostly empty: 320 cells = 2.5 tiles run at 0.77 of the copy bandwidth, 448 = 3.5 at 0.87 (profiles/r02_exp_cta_trace.txt §3).
// The 64-cell tile (two thread rows per warp, 15 core rows) covers such rows with whole tiles at the price of twice the x-halo
// columns and right-column cells and of a producer warp that also owns a row of E' (forced onto 512³, where the wide tile fits
// exactly: 119.0 against 129.7 Gcells/s); it is taken when it saves at least 10 % of the tile area — measured, wide → narrow:
// 192³ 83.3 → 94.9, 320³ 105.7 → 116.7, 448³ 117.4 → 123.1, 300³ 101.2 → 103.5, 576³ 119.4 → 120.6 (profiles/r02_exp_narrow_tiles.txt).
// FDTD_B200_NARROW_TILES=0 / 1 forces the choice.
bool use_narrow_tiles(fdtd_b200_handle h) {
    if (h->narrow_tiles < 0) {
        const long long pad128 = (static_cast<long long>(h->nx) + 127) / 128 * 128, pad64 = (static_cast<long long>(h->nx) + 63) / 64 * 64;
        bool narrow = pad64 < pad128 && static_cast<double>(pad128 - pad64) >= 0.10 * static_cast<double>(pad128);
        if (const char* e = std::getenv("FDTD_B200_NARROW_TILES")) { if (e[0] == '0') narrow = false; if (e[0] == '1') narrow = true; }
        h->narrow_tiles = narrow ? 1 : 0;
    }
    return h->narrow_tiles == 1;
}

template <int EXACT, typename IdxT, int BX, int BY>
int launch_fused_w(fdtd_b200_handle h, int z_lo, int z_hi, bool push, bool pml) {
    using LO = FusedLayout<IdxT, BX, BY>;
    constexpr int W = BX == 32 ? 0 : 1;
    if (push && pml) return fail(h, FDTD_B200_ESTATE, "the peer-memory boundary launches do not carry a PML");
    if (!h->fused_maps_ready[W]) {
        for (int set = 0; set < 2; ++set) {
            float* const* arr = (set == h->cur_set) ? h->alloc : h->alt_alloc;   // maps start at the lower ghost plane
            for (int c = 0; c < 3; ++c) {
                if (int rc = make_map(h, &h->fused_maps[W][set].e[c], CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 4, arr[c], LO::PE, LO::E_ROWS, 2)) return rc;
                if (int rc = make_map(h, &h->fused_maps[W][set].h[c], CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 4, arr[3 + c], LO::PE, LO::H_ROWS, 2)) return rc;
            }
            if (int rc = make_map(h, &h->fused_maps[W][set].mat, sizeof(IdxT) == 1 ? CU_TENSOR_MAP_DATA_TYPE_UINT8 : CU_TENSOR_MAP_DATA_TYPE_UINT16,
                                  sizeof(IdxT), h->mat, LO::PM, LO::H_ROWS, 1)) return rc;   // + the upper neighbour's first plane
        }
        h->fused_maps_ready[W] = true;
    }
    FusedView V{};
    for (int c = 0; c < 6; ++c) { V.in[c] = h->f[c]; V.out[c] = h->g[c]; }
    V.mat = h->mat; V.table = h->table; V.n_table = static_cast<int>(h->n_table);
    // The coefficient table is staged in shared memory as long as two CTAs still fit on an SM (≈ 1060 rows with 8-bit
    // indices, ≈ 1000 with 16-bit ones); larger tables are read through L1.  Measured at 512³ (tools/many_materials.py,
    // profiles/r01_many_materials.jsonl): 65–256 materials 127.4 Gcells/s from shared memory against 101–119 through L1.
    uint32_t table_smem_max = 64;
    {
        if (h->sm_smem_bytes == 0) {
            if (cudaDeviceGetAttribute(&h->sm_smem_bytes, cudaDevAttrMaxSharedMemoryPerMultiprocessor, h->device) != cudaSuccess ||
                cudaDeviceGetAttribute(&h->sm_smem_reserved, cudaDevAttrReservedSharedMemoryPerBlock, h->device) != cudaSuccess) {
                (void)cudaGetLastError();
                h->sm_smem_bytes = -1;
            }
        }
        const int sm_bytes = h->sm_smem_bytes, reserved = h->sm_smem_reserved;
        if (sm_bytes > 0) {
            const long long room = sm_bytes / 2 - reserved - static_cast<long long>(LO::smem_bytes(0, pml));
            if (room > 0) table_smem_max = std::max<uint32_t>(table_smem_max, static_cast<uint32_t>(room / static_cast<long long>(sizeof(float4))));
        }
    }
    if (h->tune.table_smem > 0) table_smem_max = static_cast<uint32_t>(h->tune.table_smem);
    V.table_in_smem = ((pml && sizeof(IdxT) == 1) || h->n_table <= table_smem_max) ? 1 : 0;   // the PML instantiation assumes it for 8-bit indices (≤ 256 rows = 4 KB)
    V.nx = static_cast<int>(h->nx); V.ny = static_cast<int>(h->ny); V.nz = static_cast<int>(h->nzl);
    V.nxl = static_cast<int>(h->nxl);
    static const bool late_issue = [] { const char* e = std::getenv("FDTD_B200_LATE_ISSUE"); return e && e[0] == '1'; }();
    V.late_issue = late_issue ? 1 : 0;
    V.z_lo = z_lo; V.z_hi = z_hi;
    V.lo_open = h->z_begin > 0 ? 1 : 0;
    V.hi_open = h->z_end < h->nz ? 1 : 0;
    for (int a = 0; a < 3; ++a) { V.dx[a] = static_cast<float>(h->cfg.dx[a]); V.inv_dx[a] = 1.0f / V.dx[a]; }
    const ForcingView F = forcing_view(h);
    const ProbeView P = probe_view(h);
    int zc = h->zc_override > 0 ? h->zc_override : h->tune.zc;
    if (zc <= 0) {
        const long long tiles = static_cast<long long>((V.nx + 4 * BX - 1) / (4 * BX)) * ((V.ny + LO::CORE_ROWS - 1) / LO::CORE_ROWS);
        // The PML instantiation marches at most 32 planes: its CTAs differ in cost (x-slab tile columns +30 %, y-slab rows and
        // z-slab planes more), neighbouring tiles drift apart over a long march and their shared halo rows fall out of L2 —
        // DRAM reads per launch at 512³, 16-cell layers: 3.84 GB with 16-plane marches, 3.96 with 24, 4.13 with 32, 4.50 with 48,
        // 4.76 with 64 (3.79 algorithmic); time: 1.446 / 1.399 / 1.397 / 1.413 / 1.517 ms (profiles/r02_exp_pml.txt).
        zc = pick_fused_zc(h, tiles, z_hi - z_lo, 1, pml ? 32 : 64);
    }
    zc = std::min(zc, 128);       // the kernel's per-plane source / probe masks and FusedPmlSmem::ZMAX cover marches of at most 128 planes (+ the one above)
    FusedPmlView Q{};
    if (pml) {
        for (int a = 0; a < 3; ++a) {
            const auto& A = h->pml[a];
            Q.n[a] = A.n; Q.map[a] = A.d_map; Q.bc_h[a] = A.d_bc[0]; Q.bc_e[a] = A.d_bc[1];
            for (int g = 0; g < 2; ++g) { Q.gh_in[a][g] = A.psi[0][g]; Q.gh_out[a][g] = A.psi_alt[g]; Q.ge[a][g] = A.psi[1][g]; }
        }
        if (V.hi_open) {
            // the plane above the slab belongs to the upper neighbour: its ψ_H of this step sits in the slot of this tick's parity
            if (!h->psi_ghost) return fail(h, FDTD_B200_ESTATE, "a fused z-slab with a PML and a neighbour above needs the peer-memory transport (fdtd_b200_p2p_*) for that neighbour's psi");
            const size_t slot = 2 * (h->ghost_plane[0] + h->ghost_plane[1] + h->ghost_plane[2]);
            const float* base = h->psi_ghost + (h->tick & 1) * slot;
            const size_t bias[3] = {static_cast<size_t>(h->nzl) * h->ny * static_cast<size_t>(h->pml[0].n),      // x: [z][y][k] at z = nzl
                                    static_cast<size_t>(h->nzl) * static_cast<size_t>(h->pml[1].n) * h->nx,      // y: [z][k][x]
                                    static_cast<size_t>(h->pml[2].n) * h->ny * h->nx};                           // z: [k][y][x] at k = n
            size_t off = 0;
            for (int a = 0; a < 3; ++a) {
                for (int g = 0; g < 2; ++g) Q.gh_ghost[a][g] = base + off + g * h->ghost_plane[a] - bias[a];
                off += 2 * h->ghost_plane[a];
            }
        }
    }
    dim3 block(BX, BY, 1);
    dim3 grid((V.nx + 4 * BX - 1) / (4 * BX), (V.ny + LO::CORE_ROWS - 1) / LO::CORE_ROWS, (z_hi - z_lo + zc - 1) / zc);
    if (pml) {
        // slab tile rows and slab chunks first (FusedPmlView::heavy): the leading and the trailing run of blocks that hold a slab coordinate
        auto runs = [](const std::vector<int>& pos, int origin, int extent, int count, int out[2]) {
            std::vector<char> hit(static_cast<size_t>(count), 0);
            for (int c : pos) { const int b = (c - origin) / extent; if (c >= origin && b < count) hit[static_cast<size_t>(b)] = 1; }
            int p = 0, s = count;
            while (p < count && hit[static_cast<size_t>(p)]) ++p;
            while (s > p && hit[static_cast<size_t>(s - 1)]) --s;
            out[0] = p; out[1] = s;
        };
        if (h->heavy_key[0] != z_lo || h->heavy_key[1] != z_hi || h->heavy_key[2] != zc) {   // once per launch shape, not per step
            runs(h->pml[1].pos, 0, LO::CORE_ROWS, static_cast<int>(grid.y), h->heavy_val[0]);
            runs(h->pml[2].pos, z_lo, zc, static_cast<int>(grid.z), h->heavy_val[1]);
            h->heavy_key[0] = z_lo; h->heavy_key[1] = z_hi; h->heavy_key[2] = zc;
        }
        Q.heavy[0][0] = 0; Q.heavy[0][1] = static_cast<int>(grid.x);
        for (int i = 0; i < 2; ++i) { Q.heavy[1][i] = h->heavy_val[0][i]; Q.heavy[2][i] = h->heavy_val[1][i]; }
    }
    const size_t smem = LO::smem_bytes(V.table_in_smem ? V.n_table : 0, pml);
    const bool xpad = h->nxl != h->nx;
    auto k = xpad ? k_step_fused<EXACT, IdxT, BX, BY, false, false, true> : k_step_fused<EXACT, IdxT, BX, BY, false, false, false>;
    if constexpr (BX == 32) {   // the CPML and the peer-memory boundary instantiations exist for the 128-cell tile only
        if (pml) k = xpad ? k_step_fused<EXACT, IdxT, BX, BY, false, true, true> : k_step_fused<EXACT, IdxT, BX, BY, false, true, false>;
        else if (push) k = xpad ? k_step_fused<EXACT, IdxT, BX, BY, true, false, true> : k_step_fused<EXACT, IdxT, BX, BY, true, false, false>;
    }
    else if (pml || push) return fail(h, FDTD_B200_ESTATE, "internal: the 64-cell tile has no CPML / boundary instantiation");
    if (push) {   // boundary planes of the peer-memory transport: results also go into the neighbours' pending-set ghost planes
        const int pend = h->cur_set ^ 1;
        if (h->lower.present)
            for (int c = 0; c < 5; ++c) V.push_down[c] = h->lower.field[pend][c] + h->plane * (h->lower.nzl + 1);
        if (h->upper.present)
            for (int c = 0; c < 2; ++c) V.push_up[c] = h->upper.field[pend][c];
    }
    if (raise_smem_limit(h, k, smem)) return fail(h, FDTD_B200_ECUDA, "cudaFuncSetAttribute(MaxDynamicSharedMemorySize) failed");
#ifdef FDTD_B200_CTA_TRACE
    {   // diagnostic build: the last launch's per-CTA stamps go to the file FDTD_B200_CTA_TRACE_FILE names when the handle is destroyed
        static unsigned long long* d_trace = nullptr;
        if (!d_trace) {
            CU_TRY(h, cudaMalloc(&d_trace, sizeof(unsigned long long) * 4 * 65536));
            CU_TRY(h, cudaMemcpyToSymbol(g_cta_trace, &d_trace, sizeof(d_trace)));
        }
        h->trace_buf = d_trace; h->trace_ctas = static_cast<int>(std::min<size_t>(65536, static_cast<size_t>(grid.x) * grid.y * grid.z));
        h->trace_grid[0] = grid.x; h->trace_grid[1] = grid.y; h->trace_grid[2] = grid.z; h->trace_grid[3] = zc;
    }
#endif
    CU_TRY(h, launch_step_kernel(k, grid, block, smem, h->stream, !(h->lower.present || h->upper.present), h->fused_maps[W][h->cur_set], V, F, P, zc, Q));
    h->launches += 1;
    CU_TRY(h, cudaGetLastError());
    return FDTD_B200_OK;
}

template <int EXACT, typename IdxT>
int launch_fused_t(fdtd_b200_handle h, int z_lo, int z_hi, bool push, bool pml) {
    if (!push && !pml && use_narrow_tiles(h)) return launch_fused_w<EXACT, IdxT, 16, 16>(h, z_lo, z_hi, push, pml);
    return launch_fused_w<EXACT, IdxT, 32, 8>(h, z_lo, z_hi, push, pml);
}

// ---- CPML passes ---------------------------------------------------------------------------------

// Does the fused kernel's PML instantiation (kernels_fused.cuh) run the recursion, or the passes over the slabs?  In the
// kernel: fused handles whose x slabs put at most FusedPmlSmem::XS cells into one 128-cell tile row (a slab may cross a tile
// boundary: the recomputed column right of a tile then advances its own x-slab ψ_H from the current set).  FDTD_B200_PML_PASSES=1 forces
// the passes (the two ways agree bit for bit: tests/test_pml_gpu.py).
void pml_refresh_mode(fdtd_b200_handle h) {
    h->resident = -1;   // a CPML changes whether the L2-resident multi-step kernel may run (decided again at the next step)
    h->pml_in_kernel = false;
    h->heavy_key[0] = h->heavy_key[1] = h->heavy_key[2] = -1;
    if (!h->has_pml || !h->fused) return;
    if (const char* e = std::getenv("FDTD_B200_PML_PASSES")) if (e[0] == '1') return;
    for (int a = 0; a < 3; ++a) {
        if (h->pml[a].n && (!h->pml[a].d_map || !h->pml[a].psi_alt[0] || !h->pml[a].psi_alt[1])) return;
        if (h->pml[a].ghost && !h->pml[a].d_map) return;
    }
    // the kernel indexes the x-slab arrays [z][y][k] with 32 bits (one more plane: the ghost above a z-slab)
    if (static_cast<unsigned long long>(h->pml[0].n) * h->ny * (static_cast<unsigned long long>(h->nzl) + 1) >= (1ull << 31)) return;
    std::vector<int> per_tile(static_cast<size_t>(h->nx / 128 + 1), 0);
    for (int x : h->pml[0].pos)
        if (++per_tile[static_cast<size_t>(x / 128)] > 32) return;
    h->pml_in_kernel = true;
}

void pml_free_axis(fdtd_b200_handle h, int a) {
    auto& A = h->pml[a];
    if (A.d_pos) cudaFree(A.d_pos);
    for (int k = 0; k < 2; ++k) if (A.d_bc[k]) cudaFree(A.d_bc[k]);
    for (int fld = 0; fld < 2; ++fld) for (int g = 0; g < 2; ++g) if (A.psi[fld][g]) cudaFree(A.psi[fld][g]);
    for (int g = 0; g < 2; ++g) if (A.psi_alt[g]) cudaFree(A.psi_alt[g]);
    if (A.d_map) cudaFree(A.d_map);
    A = fdtd_b200_solver::PmlAxis{};
    h->has_pml = h->pml[0].n || h->pml[1].n || h->pml[2].n || h->pml[2].ghost;
    pml_refresh_mode(h);
}

template <int FIELD, int AXIS, typename IdxT, int EXACT>
int launch_pml_axis(fdtd_b200_handle h, const LatticeView& L, int z_lo, int z_hi) {
    const auto& A = h->pml[AXIS];
    if (A.n == 0) return FDTD_B200_OK;
    PmlAxisView P{};
    P.n = A.n; P.pos = A.d_pos; P.bc = A.d_bc[FIELD]; P.g0 = A.psi[FIELD][0]; P.g1 = A.psi[FIELD][1]; P.k0 = 0;
    int r0, r1, r2;
    if (AXIS == 0) { r0 = A.n; r1 = L.ny; r2 = z_hi - z_lo; }
    else if (AXIS == 1) { r0 = L.nx; r1 = A.n; r2 = z_hi - z_lo; }
    else {
        const auto lo = std::lower_bound(A.pos.begin(), A.pos.end(), z_lo);
        const auto hi = std::lower_bound(A.pos.begin(), A.pos.end(), z_hi);
        P.k0 = static_cast<int>(lo - A.pos.begin());
        r0 = L.nx; r1 = L.ny; r2 = static_cast<int>(hi - lo);
    }
    if (r2 <= 0) return FDTD_B200_OK;
    const float* s0 = FIELD == 0 ? L.ex : L.hx; const float* s1 = FIELD == 0 ? L.ey : L.hy; const float* s2 = FIELD == 0 ? L.ez : L.hz;
    float* d0 = FIELD == 0 ? L.hx : L.ex; float* d1 = FIELD == 0 ? L.hy : L.ey; float* d2 = FIELD == 0 ? L.hz : L.ez;
    const dim3 block(32, 8, 1);
    for (int zz = 0; zz < r2; zz += 65535) {  // gridDim.z limit
        const int nzz = std::min(r2 - zz, 65535);
        dim3 grid((r0 + 31) / 32, (r1 + 7) / 8, nzz);
        if (AXIS == 0) {   // flattened (y, k) plane
            const int ctas = (r0 * r1 + 255) / 256;
            grid = dim3(std::min(ctas, 32768), (ctas + 32767) / 32768, nzz);
        }
        PmlAxisView Q = P;
        int zl = z_lo;
        if (AXIS == 2) Q.k0 = P.k0 + zz; else zl = z_lo + zz;
        k_pml_axis<FIELD, AXIS, IdxT, EXACT><<<grid, block, 0, h->stream>>>(L, Q, s0, s1, s2, d0, d1, d2, zl, z_hi);
        h->launches += 1;
    }
    CU_TRY(h, cudaGetLastError());
    return FDTD_B200_OK;
}

// ψ recursion + correction of planes [z_lo, z_hi) after the half-step `FIELD` (0 = H, 1 = E) ran on them.  `L` names the
// buffers: the E (H) pointers are differentiated, the H (E) pointers corrected.
template <int FIELD, typename IdxT, int EXACT>
int launch_pml_axes(fdtd_b200_handle h, const LatticeView& L, int z_lo, int z_hi) {
    int rc;
    if ((rc = launch_pml_axis<FIELD, 0, IdxT, EXACT>(h, L, z_lo, z_hi))) return rc;
    if ((rc = launch_pml_axis<FIELD, 1, IdxT, EXACT>(h, L, z_lo, z_hi))) return rc;
    return launch_pml_axis<FIELD, 2, IdxT, EXACT>(h, L, z_lo, z_hi);
}

int arithmetic_mode(fdtd_b200_handle h);

template <int FIELD>
int launch_pml_on(fdtd_b200_handle h, const LatticeView& L, int z_lo, int z_hi) {
    const int mode = arithmetic_mode(h);
    if (h->idx_bytes == 1)
        return mode == 0 ? launch_pml_axes<FIELD, uint8_t, 0>(h, L, z_lo, z_hi)
                         : (mode == 1 ? launch_pml_axes<FIELD, uint8_t, 1>(h, L, z_lo, z_hi) : launch_pml_axes<FIELD, uint8_t, 2>(h, L, z_lo, z_hi));
    return mode == 0 ? launch_pml_axes<FIELD, uint16_t, 0>(h, L, z_lo, z_hi)
                     : (mode == 1 ? launch_pml_axes<FIELD, uint16_t, 1>(h, L, z_lo, z_hi) : launch_pml_axes<FIELD, uint16_t, 2>(h, L, z_lo, z_hi));
}

int sample_probes(fdtd_b200_handle h, const LatticeView& L, int z_lo, int z_hi) {
    if (!h->n_probes_local) return FDTD_B200_OK;
    const ProbeView P = probe_view(h);
    k_sample_probes<<<(P.n + 127) / 128, 128, 0, h->stream>>>(L, P, z_lo, z_hi);
    h->launches += 1;
    CU_TRY(h, cudaGetLastError());
    return FDTD_B200_OK;
}

// two-kernel path: in place
template <int FIELD>
int launch_pml(fdtd_b200_handle h, int z_lo, int z_hi) {
    if (!h->has_pml) return FDTD_B200_OK;
    const LatticeView L = lattice_view(h);
    if (int rc = launch_pml_on<FIELD>(h, L, z_lo, z_hi)) return rc;
    // the E kernel sampled the probes before the correction: sample again
    return FIELD == 1 ? sample_probes(h, L, z_lo, z_hi) : FDTD_B200_OK;
}

int arithmetic_mode(fdtd_b200_handle h);

template <int EXACT, typename IdxT>
int launch_e_box(fdtd_b200_handle h, const LatticeView& L, float* const out[3], const CellBox& B) {
    const ForcingView F = forcing_view(h);
    const dim3 block(32, 8, 1);
    const int ctas = ((B.x1 - B.x0) * (B.y1 - B.y0) + 255) / 256;   // flattened (y, x) rectangle
    const dim3 grid(std::min(ctas, 32768), (ctas + 32767) / 32768, std::min(B.z1 - B.z0, 65535));
    k_update_e_box<EXACT, IdxT><<<grid, block, 0, h->stream>>>(L, F, out[0], out[1], out[2], B);
    h->launches += 1;
    CU_TRY(h, cudaGetLastError());
    return FDTD_B200_OK;
}

// Fused step with a PML (whole-lattice handles): the fused kernel wrote H'_0, E'_0 of the pending set without ψ.  Then
//   1. H pass of the PML: ψ_H from the old E (current set), H'(pending) corrected in the slabs;
//   2. E' recomputed out of place (old E of the current set, corrected H') wherever one of its H' operands changed:
//      coordinates p with p or p+1 in the H slabs of an axis;
//   3. E pass of the PML on the pending set; probes sampled from the pending set.
int fused_pml_passes(fdtd_b200_handle h) {
    const int nzl = static_cast<int>(h->nzl);
    LatticeView L = lattice_view(h);            // E = current set (old), H = pending set
    L.hx = h->g[3]; L.hy = h->g[4]; L.hz = h->g[5];
    if (int rc = launch_pml_on<0>(h, L, 0, nzl)) return rc;
    float* const out[3] = {h->g[0], h->g[1], h->g[2]};
    const int mode = arithmetic_mode(h);
    const int ext[3] = {L.nx, L.ny, L.nz};
    for (int a = 0; a < 3; ++a) {
        const auto& A = h->pml[a];
        std::vector<char> touched(static_cast<size_t>(ext[a]) + 1, 0);
        for (int k = 0; k < A.n; ++k) {
            if (!A.h_active[k]) continue;
            touched[A.pos[k]] = 1;
            if (A.pos[k] > 0) touched[A.pos[k] - 1] = 1;
        }
        for (int lo = 0; lo < ext[a];) {
            if (!touched[lo]) { ++lo; continue; }
            int hi = lo;
            while (hi < ext[a] && touched[hi]) ++hi;
            CellBox B{0, L.nx, 0, L.ny, 0, L.nz};
            if (a == 0) { B.x0 = lo; B.x1 = hi; } else if (a == 1) { B.y0 = lo; B.y1 = hi; } else { B.z0 = lo; B.z1 = hi; }
            int rc;
            if (h->idx_bytes == 1) rc = mode == 0 ? launch_e_box<0, uint8_t>(h, L, out, B) : (mode == 1 ? launch_e_box<1, uint8_t>(h, L, out, B) : launch_e_box<2, uint8_t>(h, L, out, B));
            else rc = mode == 0 ? launch_e_box<0, uint16_t>(h, L, out, B) : (mode == 1 ? launch_e_box<1, uint16_t>(h, L, out, B) : launch_e_box<2, uint16_t>(h, L, out, B));
            if (rc) return rc;
            lo = hi;
        }
    }
    L.ex = h->g[0]; L.ey = h->g[1]; L.ez = h->g[2];   // everything pending now
    if (int rc = launch_pml_on<1>(h, L, 0, nzl)) return rc;
    return sample_probes(h, L, 0, nzl);
}

// 0 fast | 1 exact with exact reciprocals (every dx, cast to f32, is a power of two) | 2 exact with IEEE division
int arithmetic_mode(fdtd_b200_handle h) {
    if (!(h->flags & FDTD_B200_EXACT_ARITHMETIC)) return 0;
    for (int a = 0; a < 3; ++a) {
        int e = 0;
        const float m = std::frexp(static_cast<float>(h->cfg.dx[a]), &e);
        if (m != 0.5f || e < -100 || e > 100) return 2;
    }
    return 1;
}

int prof_begin(fdtd_b200_handle h, int which, int planes) {
    if (!h->profiling) return FDTD_B200_OK;
    cudaEvent_t a, b;
    CU_TRY(h, cudaEventCreate(&a));
    CU_TRY(h, cudaEventCreate(&b));
    h->prof_events[which].emplace_back(a, b);
    h->prof_cells[which] += static_cast<uint64_t>(planes) * h->nxl * h->ny;   // lattice cells (pad cells are not work)
    CU_TRY(h, cudaEventRecord(a, h->stream));
    return FDTD_B200_OK;
}
int prof_end(fdtd_b200_handle h, int which) {
    if (!h->profiling) return FDTD_B200_OK;
    CU_TRY(h, cudaEventRecord(h->prof_events[which].back().second, h->stream));
    return FDTD_B200_OK;
}

int launch_h(fdtd_b200_handle h, int z_lo, int z_hi) {
    if (z_lo >= z_hi) return FDTD_B200_OK;
    const int mode = arithmetic_mode(h);
    if (int rc = prof_begin(h, 0, z_hi - z_lo)) return rc;
    int rc;
    if (use_bulk(h)) rc = launch_bulk_any(h, false, z_lo, z_hi, mode);
    else if (h->idx_bytes == 1) rc = mode == 0 ? launch_h_t<0, uint8_t>(h, z_lo, z_hi) : (mode == 1 ? launch_h_t<1, uint8_t>(h, z_lo, z_hi) : launch_h_t<2, uint8_t>(h, z_lo, z_hi));
    else rc = mode == 0 ? launch_h_t<0, uint16_t>(h, z_lo, z_hi) : (mode == 1 ? launch_h_t<1, uint16_t>(h, z_lo, z_hi) : launch_h_t<2, uint16_t>(h, z_lo, z_hi));
    if (rc) return rc;
    if ((rc = launch_pml<0>(h, z_lo, z_hi))) return rc;
    return prof_end(h, 0);
}
int launch_e(fdtd_b200_handle h, int z_lo, int z_hi) {
    if (z_lo >= z_hi) return FDTD_B200_OK;
    const int mode = arithmetic_mode(h);
    if (int rc = prof_begin(h, 1, z_hi - z_lo)) return rc;
    int rc;
    if (use_bulk(h)) rc = launch_bulk_any(h, true, z_lo, z_hi, mode);
    else if (h->idx_bytes == 1) rc = mode == 0 ? launch_e_t<0, uint8_t>(h, z_lo, z_hi) : (mode == 1 ? launch_e_t<1, uint8_t>(h, z_lo, z_hi) : launch_e_t<2, uint8_t>(h, z_lo, z_hi));
    else rc = mode == 0 ? launch_e_t<0, uint16_t>(h, z_lo, z_hi) : (mode == 1 ? launch_e_t<1, uint16_t>(h, z_lo, z_hi) : launch_e_t<2, uint16_t>(h, z_lo, z_hi));
    if (rc) return rc;
    if ((rc = launch_pml<1>(h, z_lo, z_hi))) return rc;
    return prof_end(h, 1);
}

void prof_clear(fdtd_b200_handle h) {
    for (int w = 0; w < 2; ++w) {
        for (auto& ev : h->prof_events[w]) { cudaEventDestroy(ev.first); cudaEventDestroy(ev.second); }
        h->prof_events[w].clear();
        h->prof_cells[w] = 0;
    }
}

// ---- two-dimensional lattices: the y-marching fused kernel (kernels_2d.cuh) ---------------------------------------------------
bool step2d_now(fdtd_b200_handle h) {
    if (h->step2d < 0) {
        h->step2d = 0;
        const char* e = std::getenv("FDTD_B200_STEP2D");
        if (h->fused && h->nz == 1 && h->ny >= 2 && h->tune.zc <= 0 && h->tune.variant < 0 && !(e && e[0] == '0')) h->step2d = 1;
    }
    return h->step2d == 1 && !h->has_pml;
}

template <int EXACT, typename IdxT>
int launch_2d_t(fdtd_b200_handle h) {
    auto k = k_step_2d<EXACT, IdxT>;
    const int table_in_smem = h->n_table <= static_cast<uint32_t>(k2dTableMax) ? 1 : 0;
    const int smem = k2d_smem_bytes(table_in_smem ? static_cast<int>(h->n_table) : 0);
    if (h->step2d_slots == 0) {
        int per_sm = 0;
        if (raise_smem_limit(h, k, smem)) return fail(h, FDTD_B200_ECUDA, "cudaFuncSetAttribute(MaxDynamicSharedMemorySize) failed");
        if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, k, 32 * k2dWarps, smem) != cudaSuccess || per_sm <= 0) { (void)cudaGetLastError(); per_sm = 4; }
        h->step2d_slots = per_sm * h->n_sm;
    }
    Step2dView V{};
    for (int c = 0; c < 6; ++c) { V.in[c] = h->f[c]; V.out[c] = h->g[c]; }
    V.mat = h->mat; V.table = h->table; V.n_table = static_cast<int>(h->n_table); V.table_in_smem = table_in_smem;
    V.nx = static_cast<int>(h->nx); V.ny = static_cast<int>(h->ny); V.nxl = static_cast<int>(h->nxl);
    for (int a = 0; a < 3; ++a) { V.dx[a] = static_cast<float>(h->cfg.dx[a]); V.inv_dx[a] = 1.0f / V.dx[a]; }
    const int n_xseg = (V.nx / 4 + k2dOwned - 1) / k2dOwned;
    // Rows per march: 8, measured.  A cost model like pick_fused_zc's (longer marches amortise the two extra rows, one full wave
    // beats 3.1 waves) predicts 48 rows for 4096×4096 — and is wrong: 4096×4096×1 runs at 122.5 Gcells/s with 8-row marches, 115 with
    // 12–16, 108 with 32, 100 with 48, 113 with 4 (profiles/r02_exp_2d_kernel.txt), with a 3- or a 4-slot ring alike.  Warps that
    // march independently for long drift apart and their concurrent accesses lose DRAM-page / L2 locality; short marches re-align
    // them every few rows, and the rows re-read at the seams (2 of 8) come from L2.  Smaller lattices are bound by the latency of
    // a march, not by bandwidth, and want more, shorter ones: the longest of 8, 6, 4, 3 rows that still makes ≥ 1.6 waves of work
    // items for the device's warp slots, else 2 rows.  Measured best (profiles/r02_exp_2d_kernel.txt): 600² 2 rows (10.0 µs per step
    // against 11.7 with 4), 800² 2, 1000² 2 ≈ 4, 1400² 4 (19.8 against 21.5 with 6 and 23.9 with 8), 1700² 6, 2048² and above 8.
    const double warp_slots = static_cast<double>(h->step2d_slots) * k2dWarps;
    int yc = 2;
    for (int c : {8, 6, 4, 3})
        if (static_cast<double>(n_xseg) * ((V.ny + c - 1) / c) >= 1.6 * warp_slots) { yc = c; break; }
    yc = std::max(1, std::min(yc, V.ny));
    if (const char* e = std::getenv("FDTD_B200_STEP2D_YC")) if (std::atoi(e) > 0) yc = std::atoi(e);
    const int items = n_xseg * ((V.ny + yc - 1) / yc);
    dim3 grid((items + k2dWarps - 1) / k2dWarps, 1, 1);
    CU_TRY(h, launch_step_kernel(k, dim3(grid), dim3(32 * k2dWarps), smem, h->stream, true, V, forcing_view(h), probe_view(h), yc));
    h->launches += 1;
    CU_TRY(h, cudaGetLastError());
    return FDTD_B200_OK;
}

int launch_2d(fdtd_b200_handle h, int mode) {
    if (h->idx_bytes == 1) return mode == 0 ? launch_2d_t<0, uint8_t>(h) : (mode == 1 ? launch_2d_t<1, uint8_t>(h) : launch_2d_t<2, uint8_t>(h));
    return mode == 0 ? launch_2d_t<0, uint16_t>(h) : (mode == 1 ? launch_2d_t<1, uint16_t>(h) : launch_2d_t<2, uint16_t>(h));
}

// planes [z_lo, z_hi) of the current set → the pending set (no swap)
int launch_fused(fdtd_b200_handle h, int z_lo, int z_hi, bool push = false) {
    if (z_lo >= z_hi) return FDTD_B200_OK;
    const int mode = arithmetic_mode(h);
    if (int rc = prof_begin(h, 0, z_hi - z_lo)) return rc;
    int rc;
    if (!push && z_lo == 0 && z_hi == 1 && h->nzl == 1 && step2d_now(h)) {
        if ((rc = launch_2d(h, mode))) return rc;
        return prof_end(h, 0);
    }
    const bool whole = z_lo == 0 && z_hi == static_cast<int>(h->nzl) && h->z_begin == 0 && h->z_end == h->nz;
    const bool pml = h->has_pml && h->pml_in_kernel;
    if (h->has_pml && !pml && !whole)
        return fail(h, FDTD_B200_ESTATE, "the PML passes after a fused step cover whole lattices only (plane ranges and z-slabs need the in-kernel PML, see fdtd_b200_pml_in_kernel)");
    if (h->idx_bytes == 1) rc = mode == 0 ? launch_fused_t<0, uint8_t>(h, z_lo, z_hi, push, pml) : (mode == 1 ? launch_fused_t<1, uint8_t>(h, z_lo, z_hi, push, pml) : launch_fused_t<2, uint8_t>(h, z_lo, z_hi, push, pml));
    else rc = mode == 0 ? launch_fused_t<0, uint16_t>(h, z_lo, z_hi, push, pml) : (mode == 1 ? launch_fused_t<1, uint16_t>(h, z_lo, z_hi, push, pml) : launch_fused_t<2, uint16_t>(h, z_lo, z_hi, push, pml));
    if (rc) return rc;
    if (h->has_pml && !pml && (rc = fused_pml_passes(h))) return rc;
    return prof_end(h, 0);
}

// the pending set becomes the current one
void commit_fused(fdtd_b200_handle h) {
    for (int c = 0; c < 6; ++c) { std::swap(h->f[c], h->g[c]); std::swap(h->alloc[c], h->alt_alloc[c]); }
    h->cur_set ^= 1;
    if (h->has_pml && h->pml_in_kernel)   // ψ_H written by the kernel's PML instantiation becomes current with the fields
        for (int a = 0; a < 3; ++a)
            for (int g = 0; g < 2; ++g) std::swap(h->pml[a].psi[0][g], h->pml[a].psi_alt[g]);
}

void probe_mirror_invalidate(fdtd_b200_handle h) {
    for (auto& t : h->pin_probe_tick) t = 0;
}

// one-step protocol: the probe row of the step just enqueued follows its kernel into pinned host memory without a host
// round trip; fdtd_b200_read_probes of exactly that tick then only waits for the stream
int probe_mirror_row(fdtd_b200_handle h, uint64_t tick) {
    const size_t row = static_cast<size_t>(h->n_probes_total) * 6;
    if (row == 0 || !h->d_probe_ring) return FDTD_B200_OK;
    if (h->h_probe_ring && h->row_in_host[tick % h->ring_steps]) {   // the kernels wrote the row into host memory themselves: an event behind them is all it takes
        const int slot = static_cast<int>(tick % fdtd_b200_solver::PIN_SLOTS);
        if (!h->pin_probe_ev[slot]) CU_TRY(h, cudaEventCreateWithFlags(&h->pin_probe_ev[slot], cudaEventDisableTiming));
        CU_TRY(h, cudaEventRecord(h->pin_probe_ev[slot], h->stream));
        h->pin_probe_tick[slot] = tick + 1;
        return FDTD_B200_OK;
    }
    if (h->pin_probe_floats < row) {
        if (h->pin_probe) { CU_TRY(h, cudaStreamSynchronize(h->stream)); cudaFreeHost(h->pin_probe); h->pin_probe = nullptr; h->pin_probe_floats = 0; }
        CU_TRY(h, cudaMallocHost(&h->pin_probe, sizeof(float) * row * fdtd_b200_solver::PIN_SLOTS));
        h->pin_probe_floats = row;
        probe_mirror_invalidate(h);
    }
    const int slot = static_cast<int>(tick % fdtd_b200_solver::PIN_SLOTS);
    if (!h->pin_probe_ev[slot]) CU_TRY(h, cudaEventCreateWithFlags(&h->pin_probe_ev[slot], cudaEventDisableTiming));
    CU_TRY(h, cudaMemcpyAsync(h->pin_probe + slot * h->pin_probe_floats, h->d_probe_ring + (tick % h->ring_steps) * row, row * sizeof(float),
                              cudaMemcpyDeviceToHost, h->stream));
    CU_TRY(h, cudaEventRecord(h->pin_probe_ev[slot], h->stream));
    h->pin_probe_tick[slot] = tick + 1;
    return FDTD_B200_OK;
}

// one-step calls route their probe row into the host ring (fdtd_b200_solver::h_probe_ring) for the duration of the call
struct HostRowScope {
    fdtd_b200_handle h;
    HostRowScope(fdtd_b200_handle h_, bool on) : h(h_) { h->probe_rows_to_host = on; }
    ~HostRowScope() { h->probe_rows_to_host = false; }
};

bool staged_io() {
    static const bool v = [] { const char* e = std::getenv("FDTD_B200_STAGED_IO"); return e && e[0] == '1'; }();
    return v;
}

void end_step(fdtd_b200_handle h) {
    if (h->pin_src_inflight >= 0) {   // the step that read the pinned forcing slot is enqueued: the slot is free behind it
        if (cudaEventRecord(h->pin_src_ev[h->pin_src_inflight], h->stream) != cudaSuccess) (void)cudaGetLastError();
        h->pin_src_inflight = -1;
    }
    if (h->n_probes_total && !h->row_in_host.empty()) h->row_in_host[h->tick % h->ring_steps] = (h->probe_rows_to_host && h->hd_probe_ring) ? 1 : 0;
    h->tick += 1;
    h->time += h->cfg.dt;  // accumulated (cpu/mod.rs:513), never tick*dt
    if (h->n_src) {
        h->src_row += 1;
        if (h->src_row >= h->src_rows) { h->n_src = 0; h->src_rows = 0; h->src_row = 0; h->src_vals_mapped = nullptr; }  // forcing is per step (cpu/mod.rs:370-377)
    }
}

// ---- L2-resident lattices: one persistent launch for n steps (kernels_resident.cuh) --------------------------------------------
// Decided at the first step (and again after anything that changes eligibility).  Eligible: a whole lattice (no slab
// neighbours), no CPML, at most n_sm × 512 × 4 cells (every thread of one CTA per SM keeps ≤ 4 cells in registers: 303 104 cells on
// 148 SMs; FDTD_B200_RESIDENT_MAX_CELLS lowers it), kernel choice not forced by flags / FDTD_B200_TUNE, FDTD_B200_RESIDENT != 0.
template <int EXACT, typename IdxT, int CPT>
int launch_resident_c(fdtd_b200_handle h, const ResidentView& V) {
    void* args[] = {const_cast<ResidentView*>(&V)};
    if (h->resident == 1) {
        if constexpr (CPT <= 2) {
            auto k = k_steps_cluster<EXACT, IdxT, CPT>;   // fields in distributed shared memory
            if (h->resident_ctas > 8) {
                static std::mutex mu; static std::map<std::pair<int, const void*>, bool> done;
                std::lock_guard<std::mutex> lock(mu);
                auto key = std::make_pair(h->device, reinterpret_cast<const void*>(k));
                if (!done[key]) { CU_TRY(h, cudaFuncSetAttribute(k, cudaFuncAttributeNonPortableClusterSizeAllowed, 1)); done[key] = true; }
            }
            cudaLaunchConfig_t cfg{};
            cfg.gridDim = dim3(h->resident_ctas, 1, 1);
            cfg.blockDim = dim3(kResidentThreads, 1, 1);
            cfg.dynamicSmemBytes = 0;
            cfg.stream = h->stream;
            cudaLaunchAttribute attr[1];
            attr[0].id = cudaLaunchAttributeClusterDimension;
            attr[0].val.clusterDim.x = h->resident_ctas; attr[0].val.clusterDim.y = 1; attr[0].val.clusterDim.z = 1;
            cfg.attrs = attr; cfg.numAttrs = 1;
            CU_TRY(h, cudaLaunchKernelEx(&cfg, k, V));
        } else {
            return fail(h, FDTD_B200_ESTATE, "the cluster regime keeps at most 2 cells per thread");
        }
    } else {
        auto k = k_steps_resident<EXACT, IdxT, false, CPT>;
        CU_TRY(h, cudaLaunchCooperativeKernel(reinterpret_cast<const void*>(k), dim3(h->resident_ctas, 1, 1), dim3(kResidentThreads, 1, 1), args, 0, h->stream));
    }
    h->launches += 1;
    return FDTD_B200_OK;
}

template <int EXACT, typename IdxT>
int launch_resident_t(fdtd_b200_handle h, const ResidentView& V) {
    switch (h->resident_cpt) {
        case 1: return launch_resident_c<EXACT, IdxT, 1>(h, V);
        case 2: return launch_resident_c<EXACT, IdxT, 2>(h, V);
        default: return launch_resident_c<EXACT, IdxT, 4>(h, V);
    }
}

// CTAs of 512 threads × cpt cells that cover n cells, for the smallest cpt in {1, 2, 4} that needs at most max_ctas; 0 if none does
int resident_shape(long long n, int max_ctas, int max_cpt, int* cpt_out) {
    for (int cpt : {1, 2, 4}) {
        if (cpt > max_cpt) break;
        const long long ctas = (n + static_cast<long long>(kResidentThreads) * cpt - 1) / (static_cast<long long>(kResidentThreads) * cpt);
        if (ctas <= max_ctas) { *cpt_out = cpt; return static_cast<int>(std::max<long long>(ctas, 1)); }
    }
    return 0;
}

void resident_decide(fdtd_b200_handle h, bool allow_cluster = true) {
    h->resident = 0;
    h->resident_ctas = 0;
    h->resident_why_not.clear();
    auto no = [&](const std::string& why) { h->resident_why_not = why; };
    if (const char* e = std::getenv("FDTD_B200_RESIDENT")) if (e[0] == '0') return no("FDTD_B200_RESIDENT=0");
    if (h->flags & FDTD_B200_FORCE_GENERIC_KERNELS) return no("FDTD_B200_FORCE_GENERIC_KERNELS");
    if (h->tune.variant >= 0 || h->tune.zc > 0 || h->tune.bx > 0) return no("FDTD_B200_TUNE forces a kernel variant");
    if (h->z_begin != 0 || h->z_end != h->nz) return no("z-slab handle");
    if (h->has_pml) return no("CPML");
    if (h->n_sm == 0) {
        if (cudaDeviceGetAttribute(&h->n_sm, cudaDevAttrMultiProcessorCount, h->device) != cudaSuccess || h->n_sm <= 0) { (void)cudaGetLastError(); h->n_sm = 148; }
    }
    const long long n = static_cast<long long>(h->n_local);
    long long max_cells = static_cast<long long>(h->n_sm) * kResidentThreads * 4;
    if (const char* e = std::getenv("FDTD_B200_RESIDENT_MAX_CELLS")) max_cells = std::min<long long>(max_cells, std::atoll(e));
    if (n > max_cells) return no("lattice state does not fit the on-chip budget (" + std::to_string(n) + " cells > " + std::to_string(max_cells) + ")");
    int cpt = 0;
    if (allow_cluster) {
        if (const int ctas = resident_shape(n, 16, 2, &cpt)) {   // one thread-block cluster: fields in distributed shared memory, hardware barrier
            h->resident = 1; h->resident_ctas = ctas; h->resident_cpt = cpt;
            return;
        }
    }
    int coop = 0;
    if (cudaDeviceGetAttribute(&coop, cudaDevAttrCooperativeLaunch, h->device) != cudaSuccess || !coop) { (void)cudaGetLastError(); return no("device does not support cooperative launches"); }
    if (const int ctas = resident_shape(n, h->n_sm, 4, &cpt)) { h->resident = 2; h->resident_ctas = ctas; h->resident_cpt = cpt; return; }
    no("lattice does not fit one CTA per SM");
}

bool resident_now(fdtd_b200_handle h) {
    if (h->resident < 0) resident_decide(h);
    return h->resident > 0;
}

void end_step(fdtd_b200_handle h);
int probe_mirror_row(fdtd_b200_handle h, uint64_t tick);

// n steps in one launch (chunks of at most 4096 steps, the default probe ring)
int step_resident(fdtd_b200_handle h, uint64_t n_steps) {
    const int mode = arithmetic_mode(h);
    uint64_t done = 0;
    while (done < n_steps) {
        int chunk = static_cast<int>(std::min<uint64_t>(n_steps - done, std::max<uint32_t>(std::min<uint32_t>(h->ring_steps, 4096u), 1u)));
        if (h->resident == 1 && h->n_probes_local) {   // the cluster kernel stages its probe rows in shared memory until the launch ends
            const long long cells_per_cta = static_cast<long long>(kResidentThreads) * h->resident_cpt;
            size_t busiest = 1, run = 0; long long cta = -1;
            for (long long c : h->probe_cells_host) {
                if (c / cells_per_cta != cta) { cta = c / cells_per_cta; run = 0; }
                busiest = std::max(busiest, ++run);
            }
            chunk = std::max(1, std::min<int>(chunk, static_cast<int>(kProbeStageFloats / (6 * busiest))));
            if (6 * busiest > static_cast<size_t>(kProbeStageFloats)) {   // more probes in one CTA than the staging area holds: grid regime
                resident_decide(h, false);
                if (h->resident != 2) return -1;
                continue;
            }
        }
        ResidentView V{};
        V.L = lattice_view(h);
        V.n_steps = chunk;
        const ForcingView F = forcing_view(h);
        V.n_src = F.n; V.src_cells = F.cells; V.src_vals = F.vals;
        V.src_rows = F.n ? static_cast<int>(std::min<uint64_t>(h->src_rows - h->src_row, static_cast<uint64_t>(chunk))) : 0;
        if (h->n_probes_local) {
            V.n_probes = static_cast<int>(h->n_probes_local);
            V.probe_cells = h->d_probe_cells; V.probe_slot = h->d_probe_slot;
            V.ring = (h->probe_rows_to_host && h->hd_probe_ring) ? h->hd_probe_ring : h->d_probe_ring;
            V.tick0 = h->tick; V.ring_steps = h->ring_steps; V.n_probes_total = h->n_probes_total;
        }
        if (h->resident == 2 && !h->d_sync) {
            CU_TRY(h, cudaMalloc(&h->d_sync, sizeof(unsigned long long) * 1024));
            CU_TRY(h, cudaMemsetAsync(h->d_sync, 0, sizeof(unsigned long long) * 1024, h->stream));
        }
        V.sync_flags = h->d_sync;
        V.epoch0 = h->sync_epoch;
        static const int barrier_mode = [] { const char* e = std::getenv("FDTD_B200_RESIDENT_BARRIER"); return e ? std::atoi(e) : 0; }();
        V.barrier_mode = barrier_mode;
        static const bool debug = [] { const char* e = std::getenv("FDTD_B200_RESIDENT_DEBUG"); return e && e[0] == '1'; }();
        long long* d_dbg = nullptr;
        if (debug) {
            CU_TRY(h, cudaMalloc(&d_dbg, sizeof(long long) * 64 * 1024));
            CU_TRY(h, cudaMemsetAsync(d_dbg, 0, sizeof(long long) * 64 * 1024, h->stream));
            V.dbg = d_dbg;
        }
        if (int rcp = prof_begin(h, 0, static_cast<int>(h->nzl) * chunk)) return rcp;   // one event pair per launch; cells = lattice × steps
        int rc;
        if (h->idx_bytes == 1) rc = mode == 0 ? launch_resident_t<0, uint8_t>(h, V) : (mode == 1 ? launch_resident_t<1, uint8_t>(h, V) : launch_resident_t<2, uint8_t>(h, V));
        else rc = mode == 0 ? launch_resident_t<0, uint16_t>(h, V) : (mode == 1 ? launch_resident_t<1, uint16_t>(h, V) : launch_resident_t<2, uint16_t>(h, V));
        if (rc) {
            if (done != 0) return rc;
            (void)cudaGetLastError();
            if (h->resident == 1) {   // the cluster shape was refused (cluster size / free SMs of a GPC): same kernel, grid barrier
                const std::string why = h->err;
                resident_decide(h, false);
                if (h->resident == 2) continue;
                h->err = why;
            }
            // e.g. the cooperative launch does not fit next to other work: the streaming kernels take over for good, and step_path says so
            h->resident = 0;
            h->resident_why_not = "resident launch refused: " + h->err;
            return -1;
        }
        if (int rcp = prof_end(h, 0)) return rcp;
        if (d_dbg) {   // diagnostic: where thread 0 of every CTA spends a step (cycles): H compute | barrier | E compute | barrier
            std::vector<long long> t(64 * 1024);
            CU_TRY(h, cudaStreamSynchronize(h->stream));
            CU_TRY(h, cudaMemcpy(t.data(), d_dbg, sizeof(long long) * t.size(), cudaMemcpyDeviceToHost));
            cudaFree(d_dbg);
            double acc[4] = {0, 0, 0, 0}; int cnt = 0;
            for (int c = 0; c < h->resident_ctas; ++c)
                for (int st = 1; st < std::min(chunk, 8) - 1; ++st) {
                    const long long* q = &t[(c * 8 + st) * 8];
                    acc[0] += q[2] - q[0]; acc[1] += q[3] - q[2]; acc[2] += q[6] - q[3]; acc[3] += q[7] - q[6]; ++cnt;
                }
            if (cnt) std::fprintf(stderr, "[fdtd_b200 resident debug] ctas=%d cpt=%d mode=%d  H phase %.0f | barrier %.0f | E phase %.0f | barrier %.0f cycles (thread 0, mean over CTAs and steps 1..6)\n",
                                  h->resident_ctas, h->resident_cpt, h->resident, acc[0] / cnt, acc[1] / cnt, acc[2] / cnt, acc[3] / cnt);
        }
        h->sync_epoch += 2ull * chunk;
        for (int s = 0; s < chunk; ++s) end_step(h);
        done += chunk;
    }
    return FDTD_B200_OK;
}

// Sticky failure of the peer-memory transport (fdtd_b200_solver::h_timeout).  Cheap: one read of host memory.
int p2p_poisoned(fdtd_b200_handle h) {
    if (h->h_timeout && *h->h_timeout)
        return fail(h, FDTD_B200_ECUDA, "timed out waiting for a neighbour's halo flag (push #" + std::to_string(*h->h_timeout) +
                                            "): ranks out of step or a neighbour died; this handle's state is invalid (destroy it)");
    return FDTD_B200_OK;
}

int ensure_stage(fdtd_b200_handle h, size_t floats) {
    if (h->stage_floats >= floats) return FDTD_B200_OK;
    if (h->d_stage) { cudaFree(h->d_stage); h->d_stage = nullptr; h->stage_floats = 0; }
    CU_TRY(h, cudaMalloc(&h->d_stage, floats * sizeof(float)));
    h->stage_floats = floats;
    return FDTD_B200_OK;
}

bool validate_cfg(const fdtd_b200_config* cfg, std::string* why) {
    if (!cfg) { *why = "config is NULL"; return false; }
    if (cfg->nx == 0 || cfg->ny == 0 || cfg->nz == 0) { *why = "lattice dimensions must be >= 1 (fdtd/mod.rs:25-29 clamps to 1)"; return false; }
    if (cfg->nx > 0x7fffffffull || cfg->ny > 0x7fffffffull || cfg->nz > 0x7fffffffull) { *why = "lattice dimension exceeds 2^31-1"; return false; }
    for (int a = 0; a < 3; ++a)
        if (!(cfg->dx[a] > 0.0) || !std::isfinite(cfg->dx[a])) { *why = "dx must be finite and > 0"; return false; }
    if (!std::isfinite(cfg->dt)) { *why = "dt must be finite"; return false; }
    const uint64_t zb = cfg->z_begin, ze = (cfg->z_begin == 0 && cfg->z_end == 0) ? cfg->nz : cfg->z_end;
    if (zb >= ze || ze > cfg->nz) { *why = "z-slab [z_begin, z_end) is empty or outside the lattice"; return false; }
    return true;
}

int create_common(const fdtd_b200_config* cfg, const fdtd_b200_coefficients_f32* table, uint32_t n_table,
                  const uint32_t* material_ids, fdtd_b200_handle* out) {
    std::string why;
    if (!out) return fail(nullptr, FDTD_B200_EINVAL, "out handle pointer is NULL");
    *out = nullptr;
    if (!validate_cfg(cfg, &why)) return fail(nullptr, FDTD_B200_EINVAL, why);
    if (!table || n_table == 0 || n_table > 65536) return fail(nullptr, FDTD_B200_EINVAL, "coefficient table must hold 1..65536 entries");

    int n_dev = 0;
    if (cudaGetDeviceCount(&n_dev) != cudaSuccess || n_dev == 0) {
        (void)cudaGetLastError();
        return fail(nullptr, FDTD_B200_ECUDA, "no CUDA device available: fdtd_b200 has no CPU fallback");
    }
    if (cfg->device < 0 || cfg->device >= n_dev) return fail(nullptr, FDTD_B200_EINVAL, "device ordinal out of range");

    auto* h = new fdtd_b200_solver();
    h->cfg = *cfg;
    h->nxl = cfg->nx; h->ny = cfg->ny; h->nz = cfg->nz;
    {
        const uint64_t ze = (cfg->z_begin == 0 && cfg->z_end == 0) ? cfg->nz : cfg->z_end;
        h->nx = row_pitch(cfg->nx, cfg->flags, n_table, static_cast<uint64_t>(cfg->nx) * cfg->ny * (ze - cfg->z_begin));
    }
    h->n_table_user = n_table;
    if (h->nx != h->nxl) { h->pad_id = n_table; n_table += 1; }
    h->z_begin = cfg->z_begin;
    h->z_end = (cfg->z_begin == 0 && cfg->z_end == 0) ? cfg->nz : cfg->z_end;
    h->nzl = h->z_end - h->z_begin;
    h->plane = h->nx * h->ny;
    h->n_local = h->plane * h->nzl;
    h->device = cfg->device;
    h->flags = cfg->flags;
    h->vec4 = (h->nx % 4 == 0) && !(cfg->flags & FDTD_B200_FORCE_GENERIC_KERNELS);
    h->idx_bytes = n_table <= 256 ? 1 : 2;
    h->n_table = n_table;
    h->ring_steps = cfg->probe_ring_steps ? cfg->probe_ring_steps : 4096;
    h->tune = parse_tune();

    auto bail = [&](int code) { std::string m = h->err; fdtd_b200_destroy(h); g_error = m; return code; };
#define CU_NEW(expr)                                                                                            \
    do {                                                                                                        \
        cudaError_t e__ = (expr);                                                                               \
        if (e__ != cudaSuccess) {                                                                               \
            (void)cudaGetLastError();                                                                           \
            h->err = std::string(#expr) + ": " + cudaGetErrorString(e__);                                       \
            return bail(e__ == cudaErrorMemoryAllocation ? FDTD_B200_ENOMEM : FDTD_B200_ECUDA);                 \
        }                                                                                                       \
    } while (0)

    CU_NEW(cudaSetDevice(h->device));
    if (cudaDeviceGetAttribute(&h->n_sm, cudaDevAttrMultiProcessorCount, h->device) != cudaSuccess || h->n_sm <= 0) { (void)cudaGetLastError(); h->n_sm = 148; }
    CU_NEW(cudaStreamCreateWithFlags(&h->own_stream, cudaStreamNonBlocking));
    h->stream = h->own_stream;

    // fields: [ghost plane][nzl planes][ghost plane] per scalar component, zero-initialised (cpu/mod.rs:314-316)
    const size_t comp_floats = static_cast<size_t>(h->plane) * (h->nzl + 2);
    for (int c = 0; c < 6; ++c) {
        CU_NEW(cudaMalloc(&h->alloc[c], comp_floats * sizeof(float)));
        CU_NEW(cudaMemsetAsync(h->alloc[c], 0, comp_floats * sizeof(float), h->stream));
        h->f[c] = h->alloc[c] + h->plane;
    }
    h->fused_why_not = fused_refusal(cfg->flags, h->tune.variant, h->nx, h->idx_bytes);
    h->fused = h->fused_why_not.empty();
    if (h->fused) {
        for (int c = 0; c < 6; ++c) {
            CU_NEW(cudaMalloc(&h->alt_alloc[c], comp_floats * sizeof(float)));
            CU_NEW(cudaMemsetAsync(h->alt_alloc[c], 0, comp_floats * sizeof(float), h->stream));
            h->g[c] = h->alt_alloc[c] + h->plane;
        }
    }
    // material index (+8 bytes slack so the last quad's vector load never leaves the allocation)
    CU_NEW(cudaMalloc(&h->mat, (h->n_local + h->plane) * h->idx_bytes + 16));
    CU_NEW(cudaMemsetAsync(h->mat, 0, (h->n_local + h->plane) * h->idx_bytes + 16, h->stream));
    if (h->nx != h->nxl) {   // pad cells (also of the plane above the slab) → the all-zero coefficient row
        const size_t rows = static_cast<size_t>(h->ny) * (h->nzl + 1);
        const int blocks = static_cast<int>(std::min<size_t>((rows * (h->nx - h->nxl) + 255) / 256, 148 * 8));
        if (h->idx_bytes == 1) k_fill_pad_ids<uint8_t><<<blocks, 256, 0, h->stream>>>(static_cast<uint8_t*>(h->mat), h->nxl, h->nx, rows, h->pad_id);
        else k_fill_pad_ids<uint16_t><<<blocks, 256, 0, h->stream>>>(static_cast<uint16_t*>(h->mat), h->nxl, h->nx, rows, h->pad_id);
        h->launches += 1;
        CU_NEW(cudaGetLastError());
    }
    if (material_ids) {
        // stream the u32 ids (rows of nxl cells) through a bounded staging buffer and narrow them on the device into the pitched rows
        const size_t n_cells = static_cast<size_t>(h->nxl) * h->ny * h->nzl;
        const size_t rows_per_chunk = std::max<size_t>(1, (size_t(1) << 24) / h->nxl);
        const size_t chunk = std::min<size_t>(n_cells, rows_per_chunk * h->nxl);
        uint32_t* d_ids = nullptr;
        CU_NEW(cudaMalloc(&d_ids, chunk * sizeof(uint32_t)));
        for (size_t off = 0; off < n_cells; off += chunk) {
            const size_t n = std::min(chunk, n_cells - off);
            for (size_t i = 0; i < n; ++i) {
                if (material_ids[off + i] >= h->n_table_user) {
                    cudaFree(d_ids);
                    h->err = "material id out of range of the palette/table";
                    return bail(FDTD_B200_EINVAL);
                }
            }
            cudaError_t e = cudaMemcpyAsync(d_ids, material_ids + off, n * sizeof(uint32_t), cudaMemcpyHostToDevice, h->stream);
            if (e == cudaSuccess) {
                const int blocks = static_cast<int>(std::min<size_t>((n + 255) / 256, 148 * 8));
                const size_t row0 = off / h->nxl;   // chunks are whole rows
                if (h->idx_bytes == 1) k_narrow_ids<uint8_t><<<blocks, 256, 0, h->stream>>>(d_ids, static_cast<uint8_t*>(h->mat) + row0 * h->nx, n, h->nxl, h->nx);
                else k_narrow_ids<uint16_t><<<blocks, 256, 0, h->stream>>>(d_ids, static_cast<uint16_t*>(h->mat) + row0 * h->nx, n, h->nxl, h->nx);
                h->launches += 1;
                e = cudaStreamSynchronize(h->stream);
            }
            if (e != cudaSuccess) { cudaFree(d_ids); h->err = std::string("material upload: ") + cudaGetErrorString(e); return bail(FDTD_B200_ECUDA); }
        }
        cudaFree(d_ids);
    }
    CU_NEW(cudaMalloc(&h->table, sizeof(float4) * n_table));
    static_assert(sizeof(fdtd_b200_coefficients_f32) == sizeof(float4), "table entry is four floats");
    CU_NEW(cudaMemsetAsync(h->table, 0, sizeof(float4) * n_table, h->stream));   // the pad row, if any, is all zeros
    CU_NEW(cudaMemcpyAsync(h->table, table, sizeof(float4) * h->n_table_user, cudaMemcpyHostToDevice, h->stream));
    CU_NEW(cudaStreamSynchronize(h->stream));
#undef CU_NEW
    *out = h;
    return FDTD_B200_OK;
}

// A cell index of the caller's lattice (x + nxl·(y + ny·z)) → the index of the same cell in the pitched arrays.
uint64_t pitched_cell(fdtd_b200_handle h, uint64_t c) {
    if (h->nx == h->nxl) return c;
    return (c / h->nxl) * h->nx + c % h->nxl;
}

// global box → local BoxView; EINVAL when empty or outside slab
int make_box(fdtd_b200_handle h, const uint64_t lo[3], const uint64_t hi[3], BoxView* B, size_t* n_cells) {
    if (!lo || !hi) return fail(h, FDTD_B200_EINVAL, "box bounds are NULL");
    uint64_t l[3], u[3];
    const uint64_t dims[3] = {h->nxl, h->ny, h->nz};
    for (int a = 0; a < 3; ++a) {
        l[a] = std::min(lo[a], dims[a]);
        u[a] = std::min(std::max(hi[a], lo[a]), dims[a]);  // util.rs:157-160: end = max(start, end)
        if (u[a] < l[a]) u[a] = l[a];
    }
    *n_cells = static_cast<size_t>((u[0] - l[0]) * (u[1] - l[1]) * (u[2] - l[2]));
    if (*n_cells == 0) { *B = BoxView{}; return FDTD_B200_OK; }
    if (l[2] < h->z_begin || u[2] > h->z_end) return fail(h, FDTD_B200_EINVAL, "box leaves this handle's z-slab");
    B->x0 = static_cast<int>(l[0]); B->y0 = static_cast<int>(l[1]); B->z0 = static_cast<int>(l[2] - h->z_begin);
    B->bx = static_cast<int>(u[0] - l[0]); B->by = static_cast<int>(u[1] - l[1]); B->bz = static_cast<int>(u[2] - l[2]);
    B->nx = static_cast<int>(h->nx); B->ny = static_cast<int>(h->ny);
    return FDTD_B200_OK;
}

}  // namespace

// =================================================================================================
extern "C" {

int fdtd_b200_abi_version(void) { return FDTD_B200_ABI_VERSION; }

int fdtd_b200_device_count(void) {
    int n = 0;
    if (cudaGetDeviceCount(&n) != cudaSuccess) { (void)cudaGetLastError(); return 0; }
    return n;
}

const char* fdtd_b200_global_error(void) { return g_error.c_str(); }

int fdtd_b200_update_coefficients_new(double dt, const fdtd_b200_physical_constants* pc, const fdtd_b200_material* m,
                                      fdtd_b200_update_coefficients* out) {
    if (!pc || !m || !out) return fail(nullptr, FDTD_B200_EINVAL, "NULL argument");
    const bool ok1 = c_or_d(m->relative_permittivity * pc->vacuum_permittivity, m->eletrical_conductivity, dt, &out->c_a, &out->c_b);
    const bool ok2 = c_or_d(m->relative_permeability * pc->vacuum_permeability, m->magnetic_conductivity, dt, &out->d_a, &out->d_b);
    if (!ok1 || !ok2) return fail(nullptr, FDTD_B200_EINVAL, "update coefficient is NaN (fdtd/util.rs:112-113 asserts)");
    return FDTD_B200_OK;
}

int fdtd_b200_index_materials(const fdtd_b200_material* materials, uint64_t n_cells, uint32_t* ids_out,
                              fdtd_b200_material* palette_out, uint32_t palette_cap, uint32_t* n_palette_out) {
    if (!materials || !ids_out || !n_palette_out) return fail(nullptr, FDTD_B200_EINVAL, "NULL argument");
    std::unordered_map<MatKey, uint32_t, MatKeyHash> seen;
    uint32_t n = 0;
    for (uint64_t i = 0; i < n_cells; ++i) {
        const MatKey k = key_of(materials[i]);
        auto it = seen.find(k);
        if (it == seen.end()) {
            it = seen.emplace(k, n).first;
            if (n < palette_cap && palette_out) palette_out[n] = materials[i];
            ++n;
        }
        ids_out[i] = it->second;
    }
    *n_palette_out = n;
    if (n > palette_cap) return fail(nullptr, FDTD_B200_EINVAL, "more distinct materials than palette_cap");
    return FDTD_B200_OK;
}

int fdtd_b200_canonicalize_material_index(const uint32_t* ids_in, uint64_t n_cells, const fdtd_b200_material* palette_in,
                                          uint32_t n_palette_in, uint32_t* ids_out, fdtd_b200_material* palette_out,
                                          uint32_t palette_cap, uint32_t* n_palette_out) {
    if (!ids_in || !palette_in || !ids_out || !n_palette_out) return fail(nullptr, FDTD_B200_EINVAL, "NULL argument");
    // merge labels that denote bit-identical materials, then renumber by first occurrence
    std::unordered_map<MatKey, uint32_t, MatKeyHash> canon_of_key;
    std::vector<uint32_t> merged(n_palette_in);
    for (uint32_t p = 0; p < n_palette_in; ++p) merged[p] = canon_of_key.emplace(key_of(palette_in[p]), p).first->second;
    std::vector<uint32_t> new_id(n_palette_in, UINT32_MAX);
    uint32_t n = 0;
    for (uint64_t i = 0; i < n_cells; ++i) {
        const uint32_t lab = ids_in[i];
        if (lab >= n_palette_in) return fail(nullptr, FDTD_B200_EINVAL, "material id out of range of the palette");
        const uint32_t rep = merged[lab];
        if (new_id[rep] == UINT32_MAX) {
            new_id[rep] = n;
            if (n < palette_cap && palette_out) palette_out[n] = palette_in[rep];
            ++n;
        }
        ids_out[i] = new_id[rep];
    }
    *n_palette_out = n;
    if (n > palette_cap) return fail(nullptr, FDTD_B200_EINVAL, "more distinct materials than palette_cap");
    return FDTD_B200_OK;
}

size_t fdtd_b200_memory_required(const fdtd_b200_config* cfg, uint32_t index_bytes) {
    std::string why;
    if (!validate_cfg(cfg, &why) || (index_bytes != 1 && index_bytes != 2)) return 0;
    const uint64_t ze = (cfg->z_begin == 0 && cfg->z_end == 0) ? cfg->nz : cfg->z_end;
    const uint64_t nzl = ze - cfg->z_begin;
    const uint64_t pitch = row_pitch(cfg->nx, cfg->flags, index_bytes == 1 ? 1 : 65535, static_cast<uint64_t>(cfg->nx) * cfg->ny * nzl);
    const uint64_t plane = pitch * cfg->ny;
    // 6 fp32 components with two ghost planes each + the material index (cf. 136 B/cell cpu/mod.rs:217-221,
    // 80 B/cell wgpu/mod.rs:129-133); the fused step keeps two sets of the fields
    const bool fused = fused_refusal(cfg->flags, parse_tune().variant, pitch, static_cast<int>(index_bytes)).empty();
    return static_cast<size_t>((fused ? 12 : 6) * sizeof(float) * plane * (nzl + 2) + plane * (nzl + 1) * index_bytes + 16);
}

int fdtd_b200_choose_step_mode(const fdtd_b200_config* cfg, uint32_t index_bytes, uint64_t memory_limit, uint32_t* flags_out,
                               uint64_t* required_out) {
    std::string why;
    if (!validate_cfg(cfg, &why)) return fail(nullptr, FDTD_B200_EINVAL, why);
    if (index_bytes != 1 && index_bytes != 2) return fail(nullptr, FDTD_B200_EINVAL, "index_bytes must be 1 or 2");
    fdtd_b200_config c = *cfg;
    c.flags &= ~static_cast<uint32_t>(FDTD_B200_IN_PLACE_STEP);
    uint64_t need = fdtd_b200_memory_required(&c, index_bytes);
    if (memory_limit && need > memory_limit) {
        c.flags |= FDTD_B200_IN_PLACE_STEP;
        need = fdtd_b200_memory_required(&c, index_bytes);
    }
    if (flags_out) *flags_out = c.flags;
    if (required_out) *required_out = need;
    if (memory_limit && need > memory_limit)
        return fail(nullptr, FDTD_B200_ENOMEM, "lattice needs " + std::to_string(need) + " bytes of device memory, the limit is " + std::to_string(memory_limit));
    return FDTD_B200_OK;
}

int fdtd_b200_create_from_table(const fdtd_b200_config* cfg, const fdtd_b200_coefficients_f32* table, uint32_t n_table,
                                const uint32_t* material_ids, fdtd_b200_handle* out) {
    return create_common(cfg, table, n_table, material_ids, out);
}

int fdtd_b200_create(const fdtd_b200_config* cfg, const fdtd_b200_physical_constants* constants,
                     const fdtd_b200_material* palette, uint32_t n_palette, const uint32_t* material_ids,
                     fdtd_b200_handle* out) {
    if (out) *out = nullptr;
    if (!cfg || !constants || !palette || n_palette == 0 || n_palette > 65536)
        return fail(nullptr, FDTD_B200_EINVAL, "palette must hold 1..65536 materials");
    std::vector<fdtd_b200_coefficients_f32> table(n_palette);
    for (uint32_t k = 0; k < n_palette; ++k) {
        fdtd_b200_update_coefficients u{};
        const int rc = fdtd_b200_update_coefficients_new(cfg->dt, constants, &palette[k], &u);
        if (rc) return rc;
        // cast once f64 → f32 (wgpu/mod.rs:670-678)
        table[k] = {static_cast<float>(u.c_a), static_cast<float>(u.c_b), static_cast<float>(u.d_a), static_cast<float>(u.d_b)};
    }
    return create_common(cfg, table.data(), n_palette, material_ids, out);
}

void fdtd_b200_destroy(fdtd_b200_handle h) {
    if (!h) return;
    cudaSetDevice(h->device);
    if (h->own_stream) cudaStreamSynchronize(h->own_stream);
#ifdef FDTD_B200_CTA_TRACE
    if (const char* path = std::getenv("FDTD_B200_CTA_TRACE_FILE")) {
        if (h->trace_buf && h->trace_ctas > 0) {
            cudaDeviceSynchronize();
            std::vector<unsigned long long> t(static_cast<size_t>(h->trace_ctas) * 4);
            if (cudaMemcpy(t.data(), h->trace_buf, t.size() * sizeof(unsigned long long), cudaMemcpyDeviceToHost) == cudaSuccess) {
                if (FILE* f = std::fopen(path, "w")) {
                    std::fprintf(f, "# grid %d %d %d zc %d\n# cta smid start_ns end_ns x0 y0 za\n", h->trace_grid[0], h->trace_grid[1], h->trace_grid[2], h->trace_grid[3]);
                    for (int i = 0; i < h->trace_ctas; ++i)
                        std::fprintf(f, "%d %llu %llu %llu %llu %llu %llu\n", i, t[4 * i], t[4 * i + 1], t[4 * i + 2], t[4 * i + 3] >> 40, (t[4 * i + 3] >> 20) & 0xfffff, t[4 * i + 3] & 0xfffff);
                    std::fclose(f);
                }
            }
        }
    }
#endif
    for (int c = 0; c < 6; ++c) if (h->alloc[c]) cudaFree(h->alloc[c]);
    for (int c = 0; c < 6; ++c) if (h->alt_alloc[c]) cudaFree(h->alt_alloc[c]);
    if (h->mat) cudaFree(h->mat);
    if (h->table) cudaFree(h->table);
    if (h->d_src_cells) cudaFree(h->d_src_cells);
    if (h->d_src_vals) cudaFree(h->d_src_vals);
    if (h->d_probe_cells) cudaFree(h->d_probe_cells);
    if (h->d_probe_slot) cudaFree(h->d_probe_slot);
    if (h->h_probe_ring) cudaFreeHost(h->h_probe_ring);
    if (h->d_probe_ring) cudaFree(h->d_probe_ring);
    if (h->pin_src) cudaFreeHost(h->pin_src);
    if (h->pin_probe) cudaFreeHost(h->pin_probe);
    for (auto& ev : h->pin_src_ev) if (ev) cudaEventDestroy(ev);
    for (auto& ev : h->pin_probe_ev) if (ev) cudaEventDestroy(ev);
    if (h->d_stage) cudaFree(h->d_stage);
    for (int a = 0; a < 3; ++a) pml_free_axis(h, a);
    if (h->psi_ghost) cudaFree(h->psi_ghost);
    if (h->bstream) cudaStreamSynchronize(h->bstream);
    for (auto* link : {&h->lower, &h->upper}) { for (void* p : link->opened) cudaIpcCloseMemHandle(p); link->opened.clear(); }
    if (h->d_flags) cudaFree(h->d_flags);
    if (h->d_sync) cudaFree(h->d_sync);
    if (h->h_timeout) cudaFreeHost(const_cast<unsigned long long*>(h->h_timeout));
    if (h->ev_main) cudaEventDestroy(h->ev_main);
    if (h->ev_bnd) cudaEventDestroy(h->ev_bnd);
    if (h->ev_push) cudaEventDestroy(h->ev_push);
    if (h->bstream) cudaStreamDestroy(h->bstream);
    prof_clear(h);
    if (h->timer_start) cudaEventDestroy(h->timer_start);
    if (h->timer_stop) cudaEventDestroy(h->timer_stop);
    if (h->own_stream) cudaStreamDestroy(h->own_stream);
    (void)cudaGetLastError();
    delete h;
}

const char* fdtd_b200_last_error(fdtd_b200_handle h) { return h ? h->err.c_str() : g_error.c_str(); }

int fdtd_b200_reset(fdtd_b200_handle h) {
    if (!h) return fail(nullptr, FDTD_B200_EINVAL, "NULL handle");
    if (int rc = set_device(h)) return rc;
    const size_t comp_floats = static_cast<size_t>(h->plane) * (h->nzl + 2);
    for (int c = 0; c < 6; ++c) CU_TRY(h, cudaMemsetAsync(h->alloc[c], 0, comp_floats * sizeof(float), h->stream));
    if (h->fused)
        for (int c = 0; c < 6; ++c) CU_TRY(h, cudaMemsetAsync(h->alt_alloc[c], 0, comp_floats * sizeof(float), h->stream));
    if (h->h_probe_ring) {
        CU_TRY(h, cudaStreamSynchronize(h->stream));
        std::memset(h->h_probe_ring, 0, sizeof(float) * 6 * h->n_probes_total * static_cast<size_t>(h->ring_steps));
    }
    std::fill(h->row_in_host.begin(), h->row_in_host.end(), 0);
    if (h->d_probe_ring)
        CU_TRY(h, cudaMemsetAsync(h->d_probe_ring, 0, sizeof(float) * 6 * h->n_probes_total * static_cast<size_t>(h->ring_steps), h->stream));
    h->src_vals_mapped = nullptr;
    for (int a = 0; a < 3; ++a)
        for (int fld = 0; fld < 2; ++fld)
            for (int g = 0; g < 2; ++g)
                if (h->pml[a].psi[fld][g]) CU_TRY(h, cudaMemsetAsync(h->pml[a].psi[fld][g], 0, h->pml[a].floats * sizeof(float), h->stream));
    for (int a = 0; a < 3; ++a)
        for (int g = 0; g < 2; ++g)
            if (h->pml[a].psi_alt[g]) CU_TRY(h, cudaMemsetAsync(h->pml[a].psi_alt[g], 0, h->pml[a].floats * sizeof(float), h->stream));
    if (h->psi_ghost)   // the upper neighbour's psi_H receive slots (both tick parities)
        CU_TRY(h, cudaMemsetAsync(h->psi_ghost, 0, std::max<size_t>(4 * (h->ghost_plane[0] + h->ghost_plane[1] + h->ghost_plane[2]), 1) * sizeof(float), h->stream));
    h->halos_dirty = true;
    h->pin_src_inflight = -1;
    h->n_src = 0; h->src_rows = 0; h->src_row = 0;
    h->tick = 0;
    h->time = 0.0;
    h->probe_first_tick = 0;
    probe_mirror_invalidate(h);
    return FDTD_B200_OK;
}

int fdtd_b200_set_stream(fdtd_b200_handle h, void* cuda_stream) {
    if (!h) return fail(nullptr, FDTD_B200_EINVAL, "NULL handle");
    if (int rc = set_device(h)) return rc;
    CU_TRY(h, cudaStreamSynchronize(h->stream));
    h->stream = static_cast<cudaStream_t>(cuda_stream);  // NULL is a valid stream: the legacy default stream
    return FDTD_B200_OK;
}

int fdtd_b200_reset_stream(fdtd_b200_handle h) {
    if (!h) return fail(nullptr, FDTD_B200_EINVAL, "NULL handle");
    if (int rc = set_device(h)) return rc;
    CU_TRY(h, cudaStreamSynchronize(h->stream));
    h->stream = h->own_stream;
    return FDTD_B200_OK;
}

int fdtd_b200_sync(fdtd_b200_handle h) {
    if (!h) return fail(nullptr, FDTD_B200_EINVAL, "NULL handle");
    if (int rc = set_device(h)) return rc;
    CU_TRY(h, cudaStreamSynchronize(h->stream));
    if (h->bstream) CU_TRY(h, cudaStreamSynchronize(h->bstream));
    return p2p_poisoned(h);
}

int fdtd_b200_set_time(fdtd_b200_handle h, uint64_t tick, double time) {
    if (!h) return fail(nullptr, FDTD_B200_EINVAL, "NULL handle");
    if (!std::isfinite(time)) return fail(h, FDTD_B200_EINVAL, "time must be finite");
    h->tick = tick;
    h->time = time;
    h->n_src = 0; h->src_rows = 0; h->src_row = 0;
    h->src_vals_mapped = nullptr; h->pin_src_inflight = -1;
    h->halos_dirty = true;   // tick parity selects the psi receive slot and the push destination
    h->probe_first_tick = tick;
    probe_mirror_invalidate(h);
    return FDTD_B200_OK;
}

// ---- CPML -------------------------------------------------------------------------------------------

int fdtd_b200_pml_coefficients(double dt, const fdtd_b200_physical_constants* pc, double sigma, double kappa, double a, const double normal[3],
                               double b_out[3], double c_out[3]) {
    if (!pc || !normal || !b_out || !c_out) return fail(nullptr, FDTD_B200_EINVAL, "NULL argument");
    const double e0 = pc->vacuum_permittivity;
    const double b = std::exp(-((sigma / (e0 * kappa) + a / e0) * dt));   // CE 7.102 (pml.rs:44-47)
    const double c = sigma * (b - 1.0) / (sigma * kappa + (kappa * kappa) * a);  // CE 7.99 (pml.rs:50)
    for (int i = 0; i < 3; ++i) { b_out[i] = b * normal[i]; c_out[i] = c * normal[i]; }
    return FDTD_B200_OK;
}

int fdtd_b200_pml_coefficients_graded(double dt, const fdtd_b200_physical_constants* pc, const fdtd_b200_graded_pml* g, double depth,
                                      const double normal[3], double b_out[3], double c_out[3]) {
    if (!g) return fail(nullptr, FDTD_B200_EINVAL, "NULL argument");
    const double g1 = std::pow(depth, g->m);            // pml.rs:80-84
    const double g2 = std::pow(1.0 - depth, g->m_a);
    return fdtd_b200_pml_coefficients(dt, pc, g->sigma_max * g1, 1.0 + (g->kappa_max - 1.0) * g1, g->a_max * g2, normal, b_out, c_out);
}

int fdtd_b200_set_pml(fdtd_b200_handle h, int axis, const double* b_e, const double* c_e, const double* b_h, const double* c_h) {
    if (!h) return fail(nullptr, FDTD_B200_EINVAL, "NULL handle");
    if (axis < 0 || axis > 2) return fail(h, FDTD_B200_EINVAL, "axis must be 0, 1 or 2");
    if ((b_e == nullptr) != (c_e == nullptr) || (b_h == nullptr) != (c_h == nullptr) || (!b_e && b_h))
        return fail(h, FDTD_B200_EINVAL, "pass b and c together (b_h/c_h optional, NULL = the E profile)");
    if (h->p2p) return fail(h, FDTD_B200_ESTATE, "set the PML before fdtd_b200_p2p_export / connect: the neighbours exchange its layout there");
    if (int rc = set_device(h)) return rc;
    CU_TRY(h, cudaStreamSynchronize(h->stream));
    pml_free_axis(h, axis);
    if (!b_e) return FDTD_B200_OK;  // cleared
    if (!b_h) { b_h = b_e; c_h = c_e; }
    const uint64_t ext = axis == 0 ? h->nxl : (axis == 1 ? h->ny : h->nz);   // the caller's profile length
    const uint64_t lo = axis == 2 ? h->z_begin : 0, hi = axis == 2 ? h->z_end : ext;
    auto& A = h->pml[axis];
    std::vector<float2> bc_h, bc_e;
    for (uint64_t i = 0; i < ext; ++i) {
        if (!std::isfinite(b_e[i]) || !std::isfinite(c_e[i]) || !std::isfinite(b_h[i]) || !std::isfinite(c_h[i]))
            return fail(h, FDTD_B200_EINVAL, "PML coefficients must be finite");
        if ((c_e[i] == 0.0 && c_h[i] == 0.0) || i < lo || i >= hi) continue;
        A.pos.push_back(static_cast<int>(i - lo));
        A.h_active.push_back(c_h[i] != 0.0 ? 1 : 0);
        bc_h.push_back(make_float2(static_cast<float>(b_h[i]), static_cast<float>(c_h[i])));
        bc_e.push_back(make_float2(static_cast<float>(b_e[i]), static_cast<float>(c_e[i])));
    }
    A.n = static_cast<int>(A.pos.size());
    // fused z-slabs recompute H' of the neighbour's plane above them: if that plane is a PML coordinate its (b, c) ride along
    // as entry n (its ψ_H arrives through the peer-memory transport)
    const bool above = axis == 2 && h->fused && hi < ext;
    A.ghost = above && (c_e[hi] != 0.0 || c_h[hi] != 0.0);
    if (A.ghost) {
        bc_h.push_back(make_float2(static_cast<float>(b_h[hi]), static_cast<float>(c_h[hi])));
        bc_e.push_back(make_float2(static_cast<float>(b_e[hi]), static_cast<float>(c_e[hi])));
    }
    if (A.n == 0 && !A.ghost) { A = fdtd_b200_solver::PmlAxis{}; return FDTD_B200_OK; }
    A.floats = static_cast<size_t>(A.n) * (axis == 0 ? h->ny * h->nzl : (axis == 1 ? h->nx * h->nzl : h->nx * h->ny));
    auto oom = [&](cudaError_t e) {
        pml_free_axis(h, axis);
        return fail(h, e == cudaErrorMemoryAllocation ? FDTD_B200_ENOMEM : FDTD_B200_ECUDA, std::string("PML allocation: ") + cudaGetErrorString(e));
    };
    cudaError_t e;
    if (A.n && (e = cudaMalloc(&A.d_pos, sizeof(int) * A.n)) != cudaSuccess) return oom(e);
    for (int k = 0; k < 2; ++k) if ((e = cudaMalloc(&A.d_bc[k], sizeof(float2) * bc_h.size())) != cudaSuccess) return oom(e);
    for (int fld = 0; fld < 2 && A.n; ++fld)
        for (int g = 0; g < 2; ++g) {
            if ((e = cudaMalloc(&A.psi[fld][g], sizeof(float) * A.floats)) != cudaSuccess) return oom(e);
            if ((e = cudaMemsetAsync(A.psi[fld][g], 0, sizeof(float) * A.floats, h->stream)) != cudaSuccess) return oom(e);
        }
    std::vector<int> map;
    if (h->fused) {   // the fused kernel's PML instantiation: ψ_H ping-pong set and the coordinate → k map (+ the plane above a slab)
        map.assign(static_cast<size_t>(axis == 0 ? h->nx : hi - lo) + (above ? 1 : 0), -1);   // axis x: one entry per cell of the pitch
        for (int k = 0; k < A.n; ++k) map[static_cast<size_t>(A.pos[k])] = k;
        if (A.ghost) map[static_cast<size_t>(hi - lo)] = A.n;
        if ((e = cudaMalloc(&A.d_map, sizeof(int) * map.size())) != cudaSuccess) return oom(e);
        if ((e = cudaMemcpyAsync(A.d_map, map.data(), sizeof(int) * map.size(), cudaMemcpyHostToDevice, h->stream)) != cudaSuccess) return oom(e);
        for (int g = 0; g < 2 && A.n; ++g) {
            if ((e = cudaMalloc(&A.psi_alt[g], sizeof(float) * A.floats)) != cudaSuccess) return oom(e);
            if ((e = cudaMemsetAsync(A.psi_alt[g], 0, sizeof(float) * A.floats, h->stream)) != cudaSuccess) return oom(e);
        }
    }
    if (A.n && (e = cudaMemcpyAsync(A.d_pos, A.pos.data(), sizeof(int) * A.n, cudaMemcpyHostToDevice, h->stream)) != cudaSuccess) return oom(e);
    if ((e = cudaMemcpyAsync(A.d_bc[0], bc_h.data(), sizeof(float2) * bc_h.size(), cudaMemcpyHostToDevice, h->stream)) != cudaSuccess) return oom(e);
    if ((e = cudaMemcpyAsync(A.d_bc[1], bc_e.data(), sizeof(float2) * bc_e.size(), cudaMemcpyHostToDevice, h->stream)) != cudaSuccess) return oom(e);
    if ((e = cudaStreamSynchronize(h->stream)) != cudaSuccess) return oom(e);   // the host vectors go out of scope
    h->has_pml = true;
    pml_refresh_mode(h);
    if (h->fused && !h->pml_in_kernel && (h->z_begin != 0 || h->z_end != h->nz)) {
        pml_free_axis(h, axis);
        return fail(h, FDTD_B200_ESTATE, "on a fused z-slab the PML must fit the fused kernel's tiling (at most 32 PML x-coordinates "
                                         "per 128-cell tile, FDTD_B200_PML_PASSES unset); use two-kernel handles otherwise");
    }
    return FDTD_B200_OK;
}

int fdtd_b200_has_pml(fdtd_b200_handle h) { return h && h->has_pml ? 1 : 0; }
int fdtd_b200_pml_in_kernel(fdtd_b200_handle h) { return h && h->has_pml && h->pml_in_kernel ? 1 : 0; }

uint64_t fdtd_b200_pml_state_floats(fdtd_b200_handle h) {
    if (!h) return 0;
    uint64_t n = 0;
    for (int a = 0; a < 3; ++a) n += 4 * static_cast<uint64_t>(h->pml[a].floats);
    return n;
}

// ψ arrays in the order axis x, y, z × (H: G_ab, G_ac; E: G_ab, G_ac)
static int pml_state_copy(fdtd_b200_handle h, float* host, bool to_host) {
    if (!h || (!host && fdtd_b200_pml_state_floats(h))) return fail(h, FDTD_B200_EINVAL, "NULL argument");
    if (int rc = set_device(h)) return rc;
    size_t off = 0;
    for (int a = 0; a < 3; ++a)
        for (int fld = 0; fld < 2; ++fld)
            for (int g = 0; g < 2; ++g) {
                const auto& A = h->pml[a];
                if (!A.floats) continue;
                if (to_host) CU_TRY(h, cudaMemcpyAsync(host + off, A.psi[fld][g], sizeof(float) * A.floats, cudaMemcpyDeviceToHost, h->stream));
                else CU_TRY(h, cudaMemcpyAsync(A.psi[fld][g], host + off, sizeof(float) * A.floats, cudaMemcpyHostToDevice, h->stream));
                off += A.floats;
            }
    CU_TRY(h, cudaStreamSynchronize(h->stream));
    if (!to_host) h->halos_dirty = true;   // the slab below holds psi_H of our first plane
    return p2p_poisoned(h);
}
int fdtd_b200_read_pml_state(fdtd_b200_handle h, float* out) { return pml_state_copy(h, out, true); }
int fdtd_b200_write_pml_state(fdtd_b200_handle h, const float* in) { return pml_state_copy(h, const_cast<float*>(in), false); }

uint64_t fdtd_b200_tick(fdtd_b200_handle h) { return h ? h->tick : 0; }
double fdtd_b200_time(fdtd_b200_handle h) { return h ? h->time : 0.0; }
int fdtd_b200_index_bytes(fdtd_b200_handle h) { return h ? h->idx_bytes : 0; }
uint64_t fdtd_b200_row_pitch(fdtd_b200_handle h) { return h ? h->nx : 0; }
int fdtd_b200_uses_vec4(fdtd_b200_handle h) { return h && h->vec4 ? 1 : 0; }
int fdtd_b200_uses_bulk(fdtd_b200_handle h) { return h && use_bulk(h) ? 1 : 0; }
int fdtd_b200_uses_fused(fdtd_b200_handle h) { return h && h->fused ? 1 : 0; }

uint64_t fdtd_b200_memory_in_use(fdtd_b200_handle h) {
    if (!h) return 0;
    const uint64_t comp = static_cast<uint64_t>(h->plane) * (h->nzl + 2) * sizeof(float);
    uint64_t b = (h->fused ? 12 : 6) * comp + (h->n_local + h->plane) * h->idx_bytes + 16 + sizeof(float4) * h->n_table;
    b += h->src_cells_cap * sizeof(long long) + h->src_vals_cap * sizeof(float) + h->stage_floats * sizeof(float);
    if (h->d_probe_ring) b += sizeof(float) * 6 * h->n_probes_total * static_cast<uint64_t>(h->ring_steps) + h->n_probes_local * (sizeof(long long) + sizeof(int));
    for (int a = 0; a < 3; ++a) {
        const auto& A = h->pml[a];
        if (!A.n) continue;
        b += A.floats * sizeof(float) * (4 + (A.psi_alt[0] ? 2 : 0)) + static_cast<uint64_t>(A.n) * (sizeof(int) + 2 * sizeof(float2));
    }
    if (h->psi_ghost) b += 4 * (h->ghost_plane[0] + h->ghost_plane[1] + h->ghost_plane[2]) * sizeof(float);
    if (h->d_flags) b += 64;
    return b;
}

int fdtd_b200_step_path(fdtd_b200_handle h, char* buf, size_t cap) {
    if (!h) return 0;
    const int mode = arithmetic_mode(h);
    const std::string arith = mode == 0 ? "fast" : (mode == 1 ? "exact" : "exact-div");
    const std::string idx = h->idx_bytes == 1 ? "u8" : "u16";
    std::string s;
    if (h->resident < 0) resident_decide(h);
    if (h->resident > 0) {
        s = std::string("resident: ") + (h->resident == 1 ? "k_steps_cluster<" : "k_steps_resident<") + arith + "," + idx + "> — " + std::to_string(h->n_local) +
            (h->resident == 1 ? " cells held in registers / distributed shared memory, n steps per launch, " : " cells held in registers / L2, n steps per launch, ") +
            (h->resident == 1 ? "one cluster of " : "cooperative grid of ") + std::to_string(h->resident_ctas) + " CTAs x 512 threads x " + std::to_string(h->resident_cpt) + " cells; plane-range / slab calls use: ";
    }
    else if (!h->resident_why_not.empty() && h->n_local <= (1u << 22)) s = "[not resident: " + h->resident_why_not + "] ";
    if (h->fused && step2d_now(h)) {
        s += "fused 2-D: k_step_2d<" + arith + "," + idx + "> (nz == 1: y-marching single pass, warp-shuffle x neighbours)";
        if (h->nx != h->nxl) s += " xpad";
    } else if (h->fused) {
        s += "fused: k_step_fused<" + arith + "," + idx + ">";
        if (h->nx != h->nxl) s += " xpad";
        if (!h->has_pml && use_narrow_tiles(h)) s += " 64-cell tiles (row pitch " + std::to_string(h->nx) + ": the last 128-cell tile column would be at most half full)";
        if (h->has_pml) s += h->pml_in_kernel ? " + CPML inside the kernel" : " + CPML passes (x slabs do not fit the in-kernel tiling, or FDTD_B200_PML_PASSES=1)";
        if (h->p2p) s += " + peer-memory halos";
    } else {
        s += std::string("two-kernel (") + (use_bulk(h) ? "bulk" : (h->vec4 ? "vec4" : "generic")) + "): k_update_{h,e}_" +
            (use_bulk(h) ? "bulk" : (h->vec4 ? "vec4" : "generic")) + "<" + arith + "," + idx + ">; not fused: " + h->fused_why_not;
        if (h->has_pml) s += " + CPML passes";
    }
    if (buf && cap) {
        const size_t n = std::min(cap - 1, s.size());
        std::memcpy(buf, s.data(), n);
        buf[n] = 0;
    }
    return static_cast<int>(s.size());
}
uint64_t fdtd_b200_kernel_launches(fdtd_b200_handle h) { return h ? h->launches : 0; }

int fdtd_b200_set_profiling(fdtd_b200_handle h, int enabled) {
    if (!h) return fail(nullptr, FDTD_B200_EINVAL, "NULL handle");
    if (int rc = set_device(h)) return rc;
    CU_TRY(h, cudaStreamSynchronize(h->stream));
    prof_clear(h);
    h->profiling = enabled != 0;
    return FDTD_B200_OK;
}

int fdtd_b200_get_profile(fdtd_b200_handle h, int which, double* total_ms, uint64_t* n_launches, uint64_t* n_cells) {
    if (!h || which < 0 || which > 1) return fail(h, FDTD_B200_EINVAL, "which must be 0 (H) or 1 (E)");
    if (int rc = set_device(h)) return rc;
    CU_TRY(h, cudaStreamSynchronize(h->stream));
    double ms = 0.0;
    for (auto& ev : h->prof_events[which]) {
        float t = 0.f;
        CU_TRY(h, cudaEventElapsedTime(&t, ev.first, ev.second));
        ms += t;
    }
    if (total_ms) *total_ms = ms;
    if (n_launches) *n_launches = h->prof_events[which].size();
    if (n_cells) *n_cells = h->prof_cells[which];
    return FDTD_B200_OK;
}

int fdtd_b200_timer_start(fdtd_b200_handle h) {
    if (!h) return fail(nullptr, FDTD_B200_EINVAL, "NULL handle");
    if (int rc = set_device(h)) return rc;
    if (!h->timer_start) { CU_TRY(h, cudaEventCreate(&h->timer_start)); CU_TRY(h, cudaEventCreate(&h->timer_stop)); }
    CU_TRY(h, cudaStreamSynchronize(h->stream));
    CU_TRY(h, cudaEventRecord(h->timer_start, h->stream));
    return FDTD_B200_OK;
}

int fdtd_b200_timer_stop(fdtd_b200_handle h, double* elapsed_ms) {
    if (!h || !elapsed_ms) return fail(h, FDTD_B200_EINVAL, "NULL argument");
    if (!h->timer_start) return fail(h, FDTD_B200_ESTATE, "timer_stop without timer_start");
    if (int rc = set_device(h)) return rc;
    CU_TRY(h, cudaEventRecord(h->timer_stop, h->stream));
    CU_TRY(h, cudaEventSynchronize(h->timer_stop));
    float t = 0.f;
    CU_TRY(h, cudaEventElapsedTime(&t, h->timer_start, h->timer_stop));
    *elapsed_ms = t;
    return FDTD_B200_OK;
}

int fdtd_b200_read_material_ids(fdtd_b200_handle h, uint32_t* ids_out) {
    if (!h || !ids_out) return fail(h, FDTD_B200_EINVAL, "NULL argument");
    if (int rc = set_device(h)) return rc;
    const size_t n_cells = static_cast<size_t>(h->nxl) * h->ny * h->nzl;   // the caller's lattice: rows of nxl ids
    const size_t rows_per_chunk = std::max<size_t>(1, (size_t(1) << 24) / h->nxl);
    const size_t chunk = std::min<size_t>(n_cells, rows_per_chunk * h->nxl);
    uint32_t* d = nullptr;
    CU_TRY(h, cudaMalloc(&d, chunk * sizeof(uint32_t)));
    for (size_t off = 0; off < n_cells; off += chunk) {
        const size_t n = std::min(chunk, n_cells - off);
        const int blocks = static_cast<int>(std::min<size_t>((n + 255) / 256, 148 * 8));
        const size_t row0 = off / h->nxl;
        if (h->idx_bytes == 1) k_widen_ids<uint8_t><<<blocks, 256, 0, h->stream>>>(static_cast<const uint8_t*>(h->mat) + row0 * h->nx, d, n, h->nxl, h->nx);
        else k_widen_ids<uint16_t><<<blocks, 256, 0, h->stream>>>(static_cast<const uint16_t*>(h->mat) + row0 * h->nx, d, n, h->nxl, h->nx);
        h->launches += 1;
        cudaError_t e = cudaMemcpyAsync(ids_out + off, d, n * sizeof(uint32_t), cudaMemcpyDeviceToHost, h->stream);
        if (e == cudaSuccess) e = cudaStreamSynchronize(h->stream);
        if (e != cudaSuccess) { cudaFree(d); return fail(h, FDTD_B200_ECUDA, std::string("read_material_ids: ") + cudaGetErrorString(e)); }
    }
    cudaFree(d);
    return FDTD_B200_OK;
}

int fdtd_b200_read_coefficient_table(fdtd_b200_handle h, fdtd_b200_coefficients_f32* out, uint32_t cap, uint32_t* n_out) {
    if (!h || !n_out) return fail(h, FDTD_B200_EINVAL, "NULL argument");
    *n_out = h->n_table_user;   // without the pad row of padded lattices
    if (!out || cap == 0) return FDTD_B200_OK;
    if (int rc = set_device(h)) return rc;
    const uint32_t n = std::min(cap, h->n_table_user);
    CU_TRY(h, cudaMemcpyAsync(out, h->table, sizeof(float4) * n, cudaMemcpyDeviceToHost, h->stream));
    CU_TRY(h, cudaStreamSynchronize(h->stream));
    return FDTD_B200_OK;
}

// ---- forcing ---------------------------------------------------------------------------------------

int fdtd_b200_set_source_table(fdtd_b200_handle h, uint32_t n_src, const uint64_t* cells, uint64_t n_steps, const float* values) {
    if (!h) return fail(nullptr, FDTD_B200_EINVAL, "NULL handle");
    if (n_src == 0 || n_steps == 0) { h->n_src = 0; h->src_rows = 0; h->src_row = 0; h->src_vals_mapped = nullptr; h->pin_src_inflight = -1; return FDTD_B200_OK; }
    if (!cells || !values) return fail(h, FDTD_B200_EINVAL, "NULL cells/values");
    const uint64_t n_global = h->nxl * h->ny * h->nz;
    for (uint32_t k = 0; k < n_src; ++k)
        if (cells[k] >= n_global) return fail(h, FDTD_B200_EINVAL, "set_forcing called with invalid point (cell index outside the lattice)");
    if (int rc = set_device(h)) return rc;

    // per-step fast path (the reference protocol: set_forcing* → finish every step, runner.rs:522-526): the same cell list as
    // the last call → the sorted/deduplicated device table is still right and only the values move, straight into a pinned
    // slot — no allocation, no hashing, no sort, no synchronisation.  The slot is read by the step's kernels through its device
    // alias (or copied in stream order with FDTD_B200_STAGED_IO=1) and reused only after the event end_step records behind them.
    if (n_steps == 1 && h->d_src_cells && !h->src_perm.empty() && h->src_cells_raw.size() == n_src &&
        std::memcmp(h->src_cells_raw.data(), cells, sizeof(uint64_t) * n_src) == 0 && h->src_vals_cap >= h->src_perm.size() * 6) {
        const uint32_t m = static_cast<uint32_t>(h->src_perm.size());
        const size_t floats = static_cast<size_t>(m) * 6;
        if (h->pin_src_floats < floats) {
            CU_TRY(h, cudaStreamSynchronize(h->stream));
            if (h->pin_src) { cudaFreeHost(h->pin_src); h->pin_src = nullptr; h->pin_src_floats = 0; }
            const size_t cap = std::max<size_t>(floats, 96);
            CU_TRY(h, cudaHostAlloc(&h->pin_src, sizeof(float) * cap * fdtd_b200_solver::PIN_SLOTS, cudaHostAllocMapped | cudaHostAllocPortable));
            h->pin_src_floats = cap;
            for (auto& ev : h->pin_src_ev)
                if (!ev) CU_TRY(h, cudaEventCreateWithFlags(&ev, cudaEventDisableTiming));
        }
        const int slot = h->pin_src_next;
        h->pin_src_next = (slot + 1) % fdtd_b200_solver::PIN_SLOTS;
        if (cudaEventQuery(h->pin_src_ev[slot]) != cudaSuccess) {   // a never-recorded event is complete
            (void)cudaGetLastError();
            CU_TRY(h, cudaEventSynchronize(h->pin_src_ev[slot]));
        }
        float* p = h->pin_src + static_cast<size_t>(slot) * h->pin_src_floats;
        for (uint32_t q = 0; q < m; ++q) std::memcpy(p + 6 * q, values + 6 * static_cast<size_t>(h->src_perm[q]), 6 * sizeof(float));
        void* dp = nullptr;
        if (!staged_io() && cudaHostGetDevicePointer(&dp, p, 0) == cudaSuccess) {
            // the step's kernels read the few values straight from the pinned slot; end_step records the slot's event behind them
            h->src_vals_mapped = static_cast<const float*>(dp);
            h->pin_src_inflight = slot;
        } else {
            (void)cudaGetLastError();
            h->src_vals_mapped = nullptr;
            CU_TRY(h, cudaMemcpyAsync(h->d_src_vals, p, sizeof(float) * floats, cudaMemcpyHostToDevice, h->stream));
            CU_TRY(h, cudaEventRecord(h->pin_src_ev[slot], h->stream));
        }
        h->n_src = m;
        h->src_rows = 1;
        h->src_row = 0;
        return FDTD_B200_OK;
    }

    const uint64_t* raw_cells = cells;
    std::vector<uint64_t> pitched;
    if (h->nx != h->nxl) {   // padded rows: the kernels look cells up by their index in the pitched arrays
        pitched.resize(n_src);
        for (uint32_t k = 0; k < n_src; ++k) pitched[k] = pitched_cell(h, cells[k]);
        cells = pitched.data();
    }

    // keep the slab's cells; for duplicates the later column wins (cpu/mod.rs:399-408); sort by local index
    const uint64_t base = h->z_begin * h->plane;
    std::unordered_map<long long, uint32_t> last;
    for (uint32_t k = 0; k < n_src; ++k) {
        if (cells[k] < base || cells[k] >= base + h->n_local + (h->z_end < h->nz ? h->plane : 0)) continue;
        last[static_cast<long long>(cells[k] - base)] = k;
    }
    std::vector<std::pair<long long, uint32_t>> kept(last.begin(), last.end());
    std::sort(kept.begin(), kept.end());
    const uint32_t m = static_cast<uint32_t>(kept.size());
    h->src_cells_raw.clear();
    h->src_perm.clear();
    if (m == 0) { h->n_src = 0; h->src_rows = 0; h->src_row = 0; h->src_vals_mapped = nullptr; h->pin_src_inflight = -1; return FDTD_B200_OK; }

    std::vector<long long> lc(m);
    std::vector<float> vals(static_cast<size_t>(n_steps) * m * 6);
    for (uint32_t q = 0; q < m; ++q) lc[q] = kept[q].first;
    for (uint64_t s = 0; s < n_steps; ++s)
        for (uint32_t q = 0; q < m; ++q)
            std::memcpy(&vals[(s * m + q) * 6], values + (s * n_src + kept[q].second) * 6, 6 * sizeof(float));

    h->src_cells_host.clear();
    h->src_vals_mapped = nullptr;
    h->pin_src_inflight = -1;
    if (h->src_cells_cap < m) {
        if (h->d_src_cells) cudaFree(h->d_src_cells);
        h->d_src_cells = nullptr; h->src_cells_cap = 0;
        CU_TRY(h, cudaMalloc(&h->d_src_cells, sizeof(long long) * m));
        h->src_cells_cap = m;
    }
    if (h->src_vals_cap < vals.size()) {
        if (h->d_src_vals) cudaFree(h->d_src_vals);
        h->d_src_vals = nullptr; h->src_vals_cap = 0;
        CU_TRY(h, cudaMalloc(&h->d_src_vals, sizeof(float) * vals.size()));
        h->src_vals_cap = vals.size();
    }
    // previously enqueued steps may still be reading the old table
    CU_TRY(h, cudaStreamSynchronize(h->stream));
    CU_TRY(h, cudaMemcpy(h->d_src_cells, lc.data(), sizeof(long long) * m, cudaMemcpyHostToDevice));
    CU_TRY(h, cudaMemcpy(h->d_src_vals, vals.data(), sizeof(float) * vals.size(), cudaMemcpyHostToDevice));
    h->src_cells_host = lc;
    h->src_cells_raw.assign(raw_cells, raw_cells + n_src);
    h->src_perm.resize(m);
    for (uint32_t q = 0; q < m; ++q) h->src_perm[q] = kept[q].second;
    h->n_src = m;
    h->src_rows = n_steps;
    h->src_row = 0;
    return FDTD_B200_OK;
}

int fdtd_b200_set_sources(fdtd_b200_handle h, uint32_t n, const fdtd_b200_source* sources) {
    if (!h) return fail(nullptr, FDTD_B200_EINVAL, "NULL handle");
    if (n == 0) return fdtd_b200_set_source_table(h, 0, nullptr, 0, nullptr);
    if (!sources) return fail(h, FDTD_B200_EINVAL, "NULL sources");
    std::vector<uint64_t> cells(n);
    std::vector<float> vals(static_cast<size_t>(n) * 6);
    for (uint32_t k = 0; k < n; ++k) {
        cells[k] = sources[k].cell;
        for (int c = 0; c < 3; ++c) { vals[6 * k + c] = sources[k].j[c]; vals[6 * k + 3 + c] = sources[k].m[c]; }
    }
    return fdtd_b200_set_source_table(h, n, cells.data(), 1, vals.data());
}

// ---- stepping --------------------------------------------------------------------------------------

int fdtd_b200_update_h(fdtd_b200_handle h, uint64_t z_lo, uint64_t z_hi) {
    if (!h) return fail(nullptr, FDTD_B200_EINVAL, "NULL handle");
    if (z_hi > h->nzl || z_lo > z_hi) return fail(h, FDTD_B200_EINVAL, "plane range outside the slab");
    if (int rc = set_device(h)) return rc;
    return launch_h(h, static_cast<int>(z_lo), static_cast<int>(z_hi));
}

int fdtd_b200_update_e(fdtd_b200_handle h, uint64_t z_lo, uint64_t z_hi) {
    if (!h) return fail(nullptr, FDTD_B200_EINVAL, "NULL handle");
    if (z_hi > h->nzl || z_lo > z_hi) return fail(h, FDTD_B200_EINVAL, "plane range outside the slab");
    if (int rc = set_device(h)) return rc;
    return launch_e(h, static_cast<int>(z_lo), static_cast<int>(z_hi));
}

int fdtd_b200_end_step(fdtd_b200_handle h) {
    if (!h) return fail(nullptr, FDTD_B200_EINVAL, "NULL handle");
    end_step(h);
    return FDTD_B200_OK;
}

int fdtd_b200_step(fdtd_b200_handle h, uint64_t n_steps) {
    if (!h) return fail(nullptr, FDTD_B200_EINVAL, "NULL handle");
    if (int rc = set_device(h)) return rc;
    const int nzl = static_cast<int>(h->nzl);
    const HostRowScope scope(h, n_steps == 1);
    if (n_steps > 0 && resident_now(h)) {
        const int rc = step_resident(h, n_steps);
        if (rc == FDTD_B200_OK) return (n_steps == 1 && h->n_probes_total) ? probe_mirror_row(h, h->tick - 1) : FDTD_B200_OK;
        if (rc > 0) return rc;
        // rc < 0: the resident launch was refused before anything ran; h->resident is now 0 and the regular kernels take over
    }
    for (uint64_t s = 0; s < n_steps; ++s) {
        if (h->fused) {
            if (int rc = launch_fused(h, 0, nzl)) return rc;
            commit_fused(h);
        } else {
            if (int rc = launch_h(h, 0, nzl)) return rc;
            if (int rc = launch_e(h, 0, nzl)) return rc;
        }
        end_step(h);
    }
    if (n_steps == 1 && h->n_probes_total) return probe_mirror_row(h, h->tick - 1);
    return FDTD_B200_OK;
}

int fdtd_b200_halo_plane(fdtd_b200_handle h, int which, void** device_ptr, uint64_t* n_floats) {
    if (!h || !device_ptr) return fail(h, FDTD_B200_EINVAL, "NULL argument");
    const size_t top = static_cast<size_t>(h->plane) * (h->nzl - 1);
    const size_t above = static_cast<size_t>(h->plane) * h->nzl;
    const bool pending = (which & FDTD_B200_HALO_PENDING) != 0;
    if (pending && !h->fused) return fail(h, FDTD_B200_ESTATE, "the pending buffer set exists only on fused-step handles");
    float* const* f = pending ? h->g : h->f;              // plane 0 of each component
    float* const* a = pending ? h->alt_alloc : h->alloc;  // the ghost plane below
    void* p = nullptr;
    uint64_t n = h->plane;
    switch (which & ~FDTD_B200_HALO_PENDING) {
        case FDTD_B200_HALO_SEND_EX_TOP: p = f[0] + top; break;
        case FDTD_B200_HALO_SEND_EY_TOP: p = f[1] + top; break;
        case FDTD_B200_HALO_RECV_EX_LOW: p = a[0]; break;
        case FDTD_B200_HALO_RECV_EY_LOW: p = a[1]; break;
        case FDTD_B200_HALO_SEND_HX_BOTTOM: p = f[3]; break;
        case FDTD_B200_HALO_SEND_HY_BOTTOM: p = f[4]; break;
        case FDTD_B200_HALO_RECV_HX_HIGH: p = f[3] + above; break;
        case FDTD_B200_HALO_RECV_HY_HIGH: p = f[4] + above; break;
        case FDTD_B200_HALO_SEND_EX_BOTTOM: p = f[0]; break;
        case FDTD_B200_HALO_SEND_EY_BOTTOM: p = f[1]; break;
        case FDTD_B200_HALO_SEND_EZ_BOTTOM: p = f[2]; break;
        case FDTD_B200_HALO_RECV_EX_HIGH: p = f[0] + above; break;
        case FDTD_B200_HALO_RECV_EY_HIGH: p = f[1] + above; break;
        case FDTD_B200_HALO_RECV_EZ_HIGH: p = f[2] + above; break;
        case FDTD_B200_HALO_SEND_MAT_BOTTOM: p = h->mat; n = (h->plane * h->idx_bytes + 3) / 4; break;
        case FDTD_B200_HALO_RECV_MAT_HIGH: p = static_cast<unsigned char*>(h->mat) + above * h->idx_bytes; n = (h->plane * h->idx_bytes + 3) / 4; break;
        default: return fail(h, FDTD_B200_EINVAL, "unknown halo plane selector");
    }
    *device_ptr = p;
    if (n_floats) *n_floats = n;
    return FDTD_B200_OK;
}

int fdtd_b200_fused_update(fdtd_b200_handle h, uint64_t z_lo, uint64_t z_hi) {
    if (!h) return fail(nullptr, FDTD_B200_EINVAL, "NULL handle");
    if (!h->fused) return fail(h, FDTD_B200_ESTATE, "handle does not run the fused step: " + h->fused_why_not);
    if (z_hi > h->nzl || z_lo > z_hi) return fail(h, FDTD_B200_EINVAL, "plane range outside the slab");
    if (int rc = set_device(h)) return rc;
    return launch_fused(h, static_cast<int>(z_lo), static_cast<int>(z_hi));
}

int fdtd_b200_fused_commit(fdtd_b200_handle h) {
    if (!h) return fail(nullptr, FDTD_B200_EINVAL, "NULL handle");
    if (!h->fused) return fail(h, FDTD_B200_ESTATE, "handle does not run the fused step: " + h->fused_why_not);
    commit_fused(h);
    return FDTD_B200_OK;
}

int fdtd_b200_halo_copy(fdtd_b200_handle dst, int which_dst, fdtd_b200_handle src, int which_src) {
    if (!dst || !src) return fail(dst, FDTD_B200_EINVAL, "NULL handle");
    if (dst->plane != src->plane) return fail(dst, FDTD_B200_EINVAL, "halo planes differ in size");
    void* pd = nullptr; void* ps = nullptr;
    uint64_t nd = 0, ns = 0;
    if (int rc = fdtd_b200_halo_plane(dst, which_dst, &pd, &nd)) return rc;
    if (int rc = fdtd_b200_halo_plane(src, which_src, &ps, &ns)) return rc;
    if (nd != ns) return fail(dst, FDTD_B200_EINVAL, "halo planes differ in size");
    // order: src stream → event → dst stream
    cudaEvent_t ev;
    CU_TRY(src, cudaSetDevice(src->device));
    CU_TRY(src, cudaEventCreateWithFlags(&ev, cudaEventDisableTiming));
    CU_TRY(src, cudaEventRecord(ev, src->stream));
    CU_TRY(dst, cudaSetDevice(dst->device));
    CU_TRY(dst, cudaStreamWaitEvent(dst->stream, ev, 0));
    const size_t bytes = static_cast<size_t>(nd) * sizeof(float);
    if (dst->device == src->device) CU_TRY(dst, cudaMemcpyAsync(pd, ps, bytes, cudaMemcpyDeviceToDevice, dst->stream));
    else CU_TRY(dst, cudaMemcpyPeerAsync(pd, dst->device, ps, src->device, bytes, dst->stream));
    // the source plane must not be overwritten before the copy has run: src waits for it
    cudaEvent_t done;
    CU_TRY(dst, cudaEventCreateWithFlags(&done, cudaEventDisableTiming));
    CU_TRY(dst, cudaEventRecord(done, dst->stream));
    CU_TRY(src, cudaSetDevice(src->device));
    CU_TRY(src, cudaStreamWaitEvent(src->stream, done, 0));
    cudaEventDestroy(ev);
    cudaEventDestroy(done);
    return FDTD_B200_OK;
}

// ---- field access ----------------------------------------------------------------------------------

int fdtd_b200_read_field(fdtd_b200_handle h, int component, const uint64_t lo[3], const uint64_t hi[3], float* out) {
    if (!h) return fail(nullptr, FDTD_B200_EINVAL, "NULL handle");
    if (component != FDTD_B200_FIELD_E && component != FDTD_B200_FIELD_H) return fail(h, FDTD_B200_EINVAL, "component must be E or H");
    BoxView B{};
    size_t n = 0;
    if (int rc = make_box(h, lo, hi, &B, &n)) return rc;
    if (n == 0) return FDTD_B200_OK;
    if (!out) return fail(h, FDTD_B200_EINVAL, "NULL output buffer");
    if (int rc = set_device(h)) return rc;
    const int c0 = component == FDTD_B200_FIELD_E ? 0 : 3;
    // stage at most 16 Mi cells at a time, whole z-planes of the box
    const size_t plane_cells = static_cast<size_t>(B.bx) * B.by;
    const int planes_per = static_cast<int>(std::max<size_t>(1, (size_t(1) << 24) / plane_cells));
    if (int rc = ensure_stage(h, std::min<size_t>(n, plane_cells * planes_per) * 3)) return rc;
    for (int z = 0; z < B.bz; z += planes_per) {
        BoxView S = B;
        S.z0 = B.z0 + z;
        S.bz = std::min(planes_per, B.bz - z);
        const size_t cells = plane_cells * S.bz;
        const int blocks = static_cast<int>(std::min<size_t>((cells + 255) / 256, 148 * 16));
        k_gather_box<<<blocks, 256, 0, h->stream>>>(h->f[c0], h->f[c0 + 1], h->f[c0 + 2], S, h->d_stage);
        h->launches += 1;
        CU_TRY(h, cudaGetLastError());
        CU_TRY(h, cudaMemcpyAsync(out + plane_cells * z * 3, h->d_stage, cells * 3 * sizeof(float), cudaMemcpyDeviceToHost, h->stream));
        CU_TRY(h, cudaStreamSynchronize(h->stream));
    }
    return p2p_poisoned(h);
}

int fdtd_b200_write_field(fdtd_b200_handle h, int component, const uint64_t lo[3], const uint64_t hi[3], const float* in) {
    if (!h) return fail(nullptr, FDTD_B200_EINVAL, "NULL handle");
    if (component != FDTD_B200_FIELD_E && component != FDTD_B200_FIELD_H) return fail(h, FDTD_B200_EINVAL, "component must be E or H");
    BoxView B{};
    size_t n = 0;
    if (int rc = make_box(h, lo, hi, &B, &n)) return rc;
    if (n == 0) return FDTD_B200_OK;
    if (!in) return fail(h, FDTD_B200_EINVAL, "NULL input buffer");
    if (int rc = set_device(h)) return rc;
    const int c0 = component == FDTD_B200_FIELD_E ? 0 : 3;
    const size_t plane_cells = static_cast<size_t>(B.bx) * B.by;
    const int planes_per = static_cast<int>(std::max<size_t>(1, (size_t(1) << 24) / plane_cells));
    if (int rc = ensure_stage(h, std::min<size_t>(n, plane_cells * planes_per) * 3)) return rc;
    for (int z = 0; z < B.bz; z += planes_per) {
        BoxView S = B;
        S.z0 = B.z0 + z;
        S.bz = std::min(planes_per, B.bz - z);
        const size_t cells = plane_cells * S.bz;
        CU_TRY(h, cudaMemcpyAsync(h->d_stage, in + plane_cells * z * 3, cells * 3 * sizeof(float), cudaMemcpyHostToDevice, h->stream));
        const int blocks = static_cast<int>(std::min<size_t>((cells + 255) / 256, 148 * 16));
        k_scatter_box<<<blocks, 256, 0, h->stream>>>(h->f[c0], h->f[c0 + 1], h->f[c0 + 2], S, h->d_stage);
        h->launches += 1;
        CU_TRY(h, cudaGetLastError());
        CU_TRY(h, cudaStreamSynchronize(h->stream));
    }
    h->halos_dirty = true;   // fields written from outside: the neighbours' ghost planes are stale until fdtd_b200_p2p_refresh
    return p2p_poisoned(h);
}

int fdtd_b200_device_field(fdtd_b200_handle h, int c, void** device_ptr) {
    if (!h || !device_ptr || c < 0 || c > 5) return fail(h, FDTD_B200_EINVAL, "bad component");
    *device_ptr = h->f[c];
    return FDTD_B200_OK;
}

// ---- probes ----------------------------------------------------------------------------------------

int fdtd_b200_register_probes(fdtd_b200_handle h, uint32_t n, const uint64_t* cells) {
    if (!h) return fail(nullptr, FDTD_B200_EINVAL, "NULL handle");
    if (int rc = set_device(h)) return rc;
    CU_TRY(h, cudaStreamSynchronize(h->stream));
    if (h->d_probe_cells) { cudaFree(h->d_probe_cells); h->d_probe_cells = nullptr; }
    if (h->d_probe_slot) { cudaFree(h->d_probe_slot); h->d_probe_slot = nullptr; }
    if (h->h_probe_ring) { cudaFreeHost(h->h_probe_ring); h->h_probe_ring = nullptr; h->hd_probe_ring = nullptr; }
    h->row_in_host.clear();
    if (h->d_probe_ring) { cudaFree(h->d_probe_ring); h->d_probe_ring = nullptr; }
    h->n_probes_total = 0; h->n_probes_local = 0;
    h->probe_cells_host.clear();
    h->probe_first_tick = h->tick;
    probe_mirror_invalidate(h);
    if (n == 0) return FDTD_B200_OK;
    if (!cells) return fail(h, FDTD_B200_EINVAL, "NULL cells");
    const uint64_t n_global = h->nxl * h->ny * h->nz;
    const uint64_t base = h->z_begin * h->plane;
    std::vector<std::pair<long long, int>> loc;
    for (uint32_t k = 0; k < n; ++k) {
        if (cells[k] >= n_global) return fail(h, FDTD_B200_EINVAL, "probe cell outside the lattice");
        const uint64_t c = pitched_cell(h, cells[k]);
        if (c >= base && c < base + h->n_local) loc.emplace_back(static_cast<long long>(c - base), static_cast<int>(k));
    }
    std::sort(loc.begin(), loc.end());
    // the fused sampler writes one slot per distinct cell; duplicates are resolved on readback
    h->n_probes_total = n;
    const size_t ring_bytes = sizeof(float) * 6 * n * static_cast<size_t>(h->ring_steps);
    if (ring_bytes <= (size_t(64) << 20) && !staged_io()) {
        void* hp = nullptr;
        void* dp = nullptr;
        if (cudaHostAlloc(&hp, ring_bytes, cudaHostAllocMapped | cudaHostAllocPortable) == cudaSuccess &&
            cudaHostGetDevicePointer(&dp, hp, 0) == cudaSuccess) {
            std::memset(hp, 0, ring_bytes);
            h->h_probe_ring = static_cast<float*>(hp);
            h->hd_probe_ring = static_cast<float*>(dp);
        } else {
            (void)cudaGetLastError();
            if (hp) cudaFreeHost(hp);
        }
    }
    h->row_in_host.assign(h->ring_steps, 0);
    CU_TRY(h, cudaMalloc(&h->d_probe_ring, ring_bytes));
    CU_TRY(h, cudaMemset(h->d_probe_ring, 0, ring_bytes));
    if (!loc.empty()) {
        std::vector<long long> lc(loc.size());
        std::vector<int> slot(loc.size());
        for (size_t q = 0; q < loc.size(); ++q) { lc[q] = loc[q].first; slot[q] = loc[q].second; }
        CU_TRY(h, cudaMalloc(&h->d_probe_cells, sizeof(long long) * lc.size()));
        CU_TRY(h, cudaMalloc(&h->d_probe_slot, sizeof(int) * slot.size()));
        CU_TRY(h, cudaMemcpy(h->d_probe_cells, lc.data(), sizeof(long long) * lc.size(), cudaMemcpyHostToDevice));
        CU_TRY(h, cudaMemcpy(h->d_probe_slot, slot.data(), sizeof(int) * slot.size(), cudaMemcpyHostToDevice));
        h->n_probes_local = static_cast<uint32_t>(loc.size());
        h->probe_cells_host = lc;
    }
    return FDTD_B200_OK;
}

int fdtd_b200_read_probes(fdtd_b200_handle h, uint64_t first_tick, uint64_t n_steps, float* out) {
    if (!h) return fail(nullptr, FDTD_B200_EINVAL, "NULL handle");
    if (n_steps == 0 || h->n_probes_total == 0) return FDTD_B200_OK;
    if (!out) return fail(h, FDTD_B200_EINVAL, "NULL output buffer");
    if (first_tick < h->probe_first_tick || first_tick + n_steps > h->tick) return fail(h, FDTD_B200_EINVAL, "probe range not sampled (yet)");
    if (h->tick - first_tick > h->ring_steps) return fail(h, FDTD_B200_EINVAL, "probe range already overwritten in the ring");
    if (int rc = set_device(h)) return rc;
    const size_t row = static_cast<size_t>(h->n_probes_total) * 6;
    const int pslot = static_cast<int>(first_tick % fdtd_b200_solver::PIN_SLOTS);
    if (n_steps == 1 && (h->pin_probe || h->h_probe_ring) && h->pin_probe_ev[pslot] && h->pin_probe_tick[pslot] == first_tick + 1) {
        // the row reached pinned memory behind its step: wait for that alone — later steps may be in flight
        // spin on the event for a while before blocking: on lattices whose step takes a few µs the wake-up latency of
        // cudaEventSynchronize (≈ 10 µs) would be most of the per-step protocol's time
        {
            cudaError_t q = cudaErrorNotReady;
            for (int spin = 0; spin < 20000 && q == cudaErrorNotReady; ++spin) q = cudaEventQuery(h->pin_probe_ev[pslot]);
            if (q == cudaErrorNotReady) { (void)cudaGetLastError(); q = cudaEventSynchronize(h->pin_probe_ev[pslot]); }
            CU_TRY(h, q);
        }
        const bool in_host = h->h_probe_ring && h->row_in_host[first_tick % h->ring_steps];
        const float* src = in_host ? h->h_probe_ring + (first_tick % h->ring_steps) * row : h->pin_probe + pslot * h->pin_probe_floats;
        std::memcpy(out, src, row * sizeof(float));
        return p2p_poisoned(h);
    }
    CU_TRY(h, cudaStreamSynchronize(h->stream));
    if (int rc = p2p_poisoned(h)) return rc;
    uint64_t done = 0;
    while (done < n_steps) {
        const uint64_t r = (first_tick + done) % h->ring_steps;
        uint64_t run = 1;   // the longest run of rows that live in the same ring
        const bool in_host = h->h_probe_ring && h->row_in_host[r];
        while (run < n_steps - done && r + run < h->ring_steps && (h->h_probe_ring && h->row_in_host[r + run]) == in_host) ++run;
        if (in_host) std::memcpy(out + done * row, h->h_probe_ring + r * row, run * row * sizeof(float));
        else CU_TRY(h, cudaMemcpy(out + done * row, h->d_probe_ring + r * row, run * row * sizeof(float), cudaMemcpyDeviceToHost));
        done += run;
    }
    return FDTD_B200_OK;
}

// ---- projection ------------------------------------------------------------------------------------

int fdtd_b200_project(fdtd_b200_handle h, int component, const float projection[16], const float color_map[16],
                      int colormap_mode, uint32_t w, uint32_t hgt, uint8_t* out_rgba, int rgba_is_device) {
    if (!h) return fail(nullptr, FDTD_B200_EINVAL, "NULL handle");
    if (component != FDTD_B200_FIELD_E && component != FDTD_B200_FIELD_H) return fail(h, FDTD_B200_EINVAL, "component must be E or H");
    if (!projection || !color_map || !out_rgba || w == 0 || hgt == 0) return fail(h, FDTD_B200_EINVAL, "bad projection arguments");
    if (colormap_mode < 0 || colormap_mode > 2) return fail(h, FDTD_B200_EINVAL, "unknown colormap mode");
    if (int rc = set_device(h)) return rc;
    const int c0 = component == FDTD_B200_FIELD_E ? 0 : 3;
    ProjectView V{};
    V.fx = h->f[c0]; V.fy = h->f[c0 + 1]; V.fz = h->f[c0 + 2];
    V.nx = static_cast<int>(h->nx); V.ny = static_cast<int>(h->ny); V.nz_global = static_cast<int>(h->nz);
    V.nxl = static_cast<int>(h->nxl);
    V.z_begin = static_cast<int>(h->z_begin); V.nz_local = static_cast<int>(h->nzl);
    std::memcpy(V.proj, projection, sizeof V.proj);
    std::memcpy(V.cmap, color_map, sizeof V.cmap);
    V.mode = colormap_mode; V.w = static_cast<int>(w); V.h = static_cast<int>(hgt);
    const size_t bytes = static_cast<size_t>(w) * hgt * 4;
    uchar4* d_out = reinterpret_cast<uchar4*>(out_rgba);
    if (!rgba_is_device) {
        if (int rc = ensure_stage(h, (bytes + 3) / 4)) return rc;
        d_out = reinterpret_cast<uchar4*>(h->d_stage);
        CU_TRY(h, cudaMemcpyAsync(d_out, out_rgba, bytes, cudaMemcpyHostToDevice, h->stream));  // untouched pixels keep caller content
    }
    dim3 block(32, 8, 1), grid((w + 31) / 32, (hgt + 7) / 8, 1);
    k_project_rgba8<<<grid, block, 0, h->stream>>>(V, d_out);
    h->launches += 1;
    CU_TRY(h, cudaGetLastError());
    if (!rgba_is_device) {
        CU_TRY(h, cudaMemcpyAsync(out_rgba, d_out, bytes, cudaMemcpyDeviceToHost, h->stream));
        CU_TRY(h, cudaStreamSynchronize(h->stream));
    }
    return p2p_poisoned(h);
}

}  // extern "C"

// ---- scene → material lattice (kernels_scene.cuh) --------------------------------------------------------------------
namespace {

static_assert(sizeof(fdtd_b200_shape) == sizeof(SceneShape), "fdtd_b200_shape mirrors SceneShape");

// Conservative lattice-space bounds of every shape: its bounding sphere pushed through the inverse affine map, padded.
void scene_boxes(const double m[16], const uint64_t dims[3], const fdtd_b200_shape* shapes, uint32_t n, std::vector<SceneBox>* out) {
    const double a[3][3] = {{m[0], m[1], m[2]}, {m[4], m[5], m[6]}, {m[8], m[9], m[10]}};
    const double det = a[0][0] * (a[1][1] * a[2][2] - a[1][2] * a[2][1]) - a[0][1] * (a[1][0] * a[2][2] - a[1][2] * a[2][0]) +
                       a[0][2] * (a[1][0] * a[2][1] - a[1][1] * a[2][0]);
    const bool invertible = std::isfinite(det) && std::fabs(det) > 1e-300;
    double inv[3][3] = {};
    if (invertible) {
        for (int r = 0; r < 3; ++r)
            for (int c = 0; c < 3; ++c) {
                const int r1 = (c + 1) % 3, r2 = (c + 2) % 3, c1 = (r + 1) % 3, c2 = (r + 2) % 3;
                inv[r][c] = (a[r1][c1] * a[r2][c2] - a[r1][c2] * a[r2][c1]) / det;   // adjugate (cyclic cofactors)
            }
    }
    out->resize(n);
    for (uint32_t i = 0; i < n; ++i) {
        SceneBox b;
        for (int ax = 0; ax < 3; ++ax) { b.lo[ax] = 0; b.hi[ax] = static_cast<int>(dims[ax]); }
        const fdtd_b200_shape& sh = shapes[i];
        double radius = 0.0;
        if (sh.kind == FDTD_B200_SHAPE_BALL) radius = std::fabs(sh.params[0]);
        else if (sh.kind == FDTD_B200_SHAPE_CUBOID) radius = std::sqrt(double(sh.params[0]) * sh.params[0] + double(sh.params[1]) * sh.params[1] + double(sh.params[2]) * sh.params[2]);
        else radius = std::sqrt(double(sh.params[0]) * sh.params[0] + double(sh.params[1]) * sh.params[1]);
        if (invertible && std::isfinite(radius)) {
            const double d[3] = {sh.translation[0] - m[3], sh.translation[1] - m[7], sh.translation[2] - m[11]};
            for (int ax = 0; ax < 3; ++ax) {
                const double c = inv[ax][0] * d[0] + inv[ax][1] * d[1] + inv[ax][2] * d[2];
                const double ext = radius * std::sqrt(inv[ax][0] * inv[ax][0] + inv[ax][1] * inv[ax][1] + inv[ax][2] * inv[ax][2]);
                const double pad = 2.0 + 1e-4 * (std::fabs(c) + ext);      // f32 rounding of world points and of the test itself
                const double lo = std::floor(c - ext - pad), hi = std::ceil(c + ext + pad) + 1.0;
                if (std::isfinite(lo) && std::isfinite(hi)) {
                    b.lo[ax] = static_cast<int>(std::min(std::max(lo, 0.0), double(dims[ax])));
                    b.hi[ax] = static_cast<int>(std::min(std::max(hi, 0.0), double(dims[ax])));
                }
            }
        }
        (*out)[i] = b;
    }
}

struct SceneDevice {
    SceneShape* shapes = nullptr;
    SceneBox* boxes = nullptr;
    ~SceneDevice() { cudaFree(shapes); cudaFree(boxes); }
};

int scene_validate(const uint64_t dims[3], const double m[16], const fdtd_b200_shape* shapes, uint32_t n, std::string* why) {
    if (!dims || !m) { *why = "NULL argument"; return FDTD_B200_EINVAL; }
    if (n && !shapes) { *why = "shapes is NULL"; return FDTD_B200_EINVAL; }
    if (n > 1024) { *why = "at most 1024 shapes per scene (they are staged in shared memory)"; return FDTD_B200_EINVAL; }
    for (int i = 0; i < 12; ++i) if (!std::isfinite(m[i])) { *why = "solver_to_world is not finite"; return FDTD_B200_EINVAL; }
    if (m[12] != 0.0 || m[13] != 0.0 || m[14] != 0.0 || m[15] != 1.0) { *why = "solver_to_world must be affine (last row 0 0 0 1)"; return FDTD_B200_EINVAL; }
    for (uint32_t i = 0; i < n; ++i) if (shapes[i].kind > FDTD_B200_SHAPE_CYLINDER) { *why = "unknown shape kind"; return FDTD_B200_EINVAL; }
    return FDTD_B200_OK;
}

int scene_upload(fdtd_b200_handle h, const uint64_t dims[3], const double m[16], const fdtd_b200_shape* shapes, uint32_t n, cudaStream_t st,
                 SceneDevice* dev, SceneView* V) {
    std::vector<SceneBox> boxes;
    scene_boxes(m, dims, shapes, n, &boxes);
    if (n) {
        CU_TRY(h, cudaMalloc(&dev->shapes, n * sizeof(SceneShape)));
        CU_TRY(h, cudaMalloc(&dev->boxes, n * sizeof(SceneBox)));
        CU_TRY(h, cudaMemcpyAsync(dev->shapes, shapes, n * sizeof(SceneShape), cudaMemcpyHostToDevice, st));
        CU_TRY(h, cudaMemcpyAsync(dev->boxes, boxes.data(), n * sizeof(SceneBox), cudaMemcpyHostToDevice, st));
        CU_TRY(h, cudaStreamSynchronize(st));   // `boxes` is a local
    }
    for (int i = 0; i < 12; ++i) V->m[i] = m[i];
    V->shapes = dev->shapes; V->boxes = dev->boxes; V->n_shapes = static_cast<int>(n);
    V->nx = static_cast<int>(dims[0]); V->ny = static_cast<int>(dims[1]);
    V->pitch = V->nx;
    return FDTD_B200_OK;
}

template <typename OutT>
int scene_launch(fdtd_b200_handle h, const SceneView& V, OutT* out, cudaStream_t st) {
    const long long quads = static_cast<long long>((V.nx + 3) / 4) * V.ny * V.n_planes;
    if (quads <= 0) return FDTD_B200_OK;
    const int blocks = static_cast<int>(std::min<long long>((quads + 255) / 256, 148LL * 32));
    const size_t smem = static_cast<size_t>(V.n_shapes) * (sizeof(SceneShape) + sizeof(SceneBox));
    auto k = k_rasterize_scene<OutT>;
    if (smem > 48 * 1024) if (raise_smem_limit(h, k, smem)) return fail(h, FDTD_B200_ECUDA, "cudaFuncSetAttribute(MaxDynamicSharedMemorySize) failed");
    k<<<blocks, 256, smem, st>>>(V, out);
    CU_TRY(h, cudaGetLastError());
    return FDTD_B200_OK;
}

}  // namespace

extern "C" {

int fdtd_b200_rasterize_scene(int device, const uint64_t dims[3], uint64_t z_begin, uint64_t z_end, const double solver_to_world[16],
                              const fdtd_b200_shape* shapes, uint32_t n_shapes, uint32_t default_label, uint32_t* ids_out) {
    std::string why;
    if (int rc = scene_validate(dims, solver_to_world, shapes, n_shapes, &why)) return fail(nullptr, rc, why);
    if (!ids_out) return fail(nullptr, FDTD_B200_EINVAL, "ids_out is NULL");
    if (dims[0] == 0 || dims[1] == 0 || dims[2] == 0 || dims[0] > 0x7fffffffull || dims[1] > 0x7fffffffull || dims[2] > 0x7fffffffull)
        return fail(nullptr, FDTD_B200_EINVAL, "lattice dimensions out of range");
    if (z_begin >= z_end || z_end > dims[2]) return fail(nullptr, FDTD_B200_EINVAL, "plane range [z_begin, z_end) is empty or outside the lattice");
    int n_dev = 0;
    if (cudaGetDeviceCount(&n_dev) != cudaSuccess || n_dev == 0) {
        (void)cudaGetLastError();
        return fail(nullptr, FDTD_B200_ECUDA, "no CUDA device available: fdtd_b200 has no CPU fallback");
    }
    if (device < 0 || device >= n_dev) return fail(nullptr, FDTD_B200_EINVAL, "device ordinal out of range");
    CU_TRY(nullptr, cudaSetDevice(device));
    SceneDevice dev;
    SceneView V{};
    if (int rc = scene_upload(nullptr, dims, solver_to_world, shapes, n_shapes, nullptr, &dev, &V)) return rc;
    V.default_label = default_label;
    const uint64_t plane = dims[0] * dims[1];
    const uint64_t chunk_planes = std::max<uint64_t>(1, std::min<uint64_t>(z_end - z_begin, (uint64_t(1) << 26) / plane));  // ≤ 256 MB staging
    uint32_t* d_out = nullptr;
    CU_TRY(nullptr, cudaMalloc(&d_out, chunk_planes * plane * sizeof(uint32_t)));
    int rc = FDTD_B200_OK;
    for (uint64_t z = z_begin; z < z_end && rc == FDTD_B200_OK; z += chunk_planes) {
        const uint64_t np = std::min(chunk_planes, z_end - z);
        V.z0 = static_cast<long long>(z); V.n_planes = static_cast<long long>(np);
        rc = scene_launch<uint32_t>(nullptr, V, d_out, nullptr);
        if (rc == FDTD_B200_OK) {
            const cudaError_t e = cudaMemcpy(ids_out + (z - z_begin) * plane, d_out, np * plane * sizeof(uint32_t), cudaMemcpyDeviceToHost);
            if (e != cudaSuccess) { (void)cudaGetLastError(); rc = fail(nullptr, FDTD_B200_ECUDA, std::string("rasterize readback: ") + cudaGetErrorString(e)); }
        }
    }
    cudaFree(d_out);
    return rc;
}

int fdtd_b200_set_materials_from_scene(fdtd_b200_handle h, const double solver_to_world[16], const fdtd_b200_shape* shapes,
                                       uint32_t n_shapes, uint32_t default_label) {
    if (!h) return fail(nullptr, FDTD_B200_EINVAL, "handle is NULL");
    const uint64_t dims[3] = {h->nxl, h->ny, h->nz};
    std::string why;
    if (int rc = scene_validate(dims, solver_to_world, shapes, n_shapes, &why)) return fail(h, rc, why);
    if (default_label >= h->n_table_user) return fail(h, FDTD_B200_EINVAL, "default label out of range of the palette/table");
    for (uint32_t i = 0; i < n_shapes; ++i)
        if (shapes[i].label >= h->n_table_user) return fail(h, FDTD_B200_EINVAL, "shape label out of range of the palette/table");
    CU_TRY(h, cudaSetDevice(h->device));
    SceneDevice dev;
    SceneView V{};
    if (int rc = scene_upload(h, dims, solver_to_world, shapes, n_shapes, h->stream, &dev, &V)) return rc;
    V.default_label = default_label;
    V.pitch = static_cast<int>(h->nx);   // pad cells keep the pad row's id
    V.z0 = static_cast<long long>(h->z_begin);
    // the index buffer keeps one plane above the slab (the fused step recomputes H' there); fill it when it exists globally
    V.n_planes = static_cast<long long>(h->nzl + (h->z_end < h->nz ? 1 : 0));
    int rc = h->idx_bytes == 1 ? scene_launch<uint8_t>(h, V, static_cast<uint8_t*>(h->mat), h->stream)
                               : scene_launch<uint16_t>(h, V, static_cast<uint16_t*>(h->mat), h->stream);
    if (rc) return rc;
    h->launches += 1;
    CU_TRY(h, cudaStreamSynchronize(h->stream));   // the device copies of shapes/boxes die with `dev`
    return FDTD_B200_OK;
}

}  // extern "C"

// ---- NVLink peer-memory halo transport ------------------------------------------------------------------------------------
// SURVEY §8e "primary" transport: boundary planes are written into the neighbour's ghost planes through peer-mapped memory and
// ordered by a release/acquire flag that lives in the receiver's HBM; nothing on the host synchronises per step.
#include <unistd.h>

namespace {

struct P2PBlob {
    uint32_t magic, version;
    int32_t device;
    uint32_t idx_bytes;
    uint64_t pid;
    uint64_t nx, ny, nz, z_begin, z_end;
    cudaIpcMemHandle_t field[2][6], mat, flags;
    uint64_t raw_field[2][6], raw_mat, raw_flags;
    // fused PML on z-slabs: this slab's receive buffer for ψ_H of the plane above it, and the layout both sides must agree on
    cudaIpcMemHandle_t psi_ghost;
    uint64_t raw_psi_ghost;
    uint64_t ghost_plane[3];      // floats per ψ array of that plane, axis x, y, z (0: no PML there / no neighbour above)
    uint64_t first_plane[3];      // … of THIS slab's first plane: what it will push down
};
static_assert(sizeof(P2PBlob) <= FDTD_B200_P2P_BLOB_BYTES, "blob must fit the size the header promises");
constexpr uint32_t P2P_MAGIC = 0xFD7DB200u;

__global__ void k_wait_flag(const volatile unsigned long long* flag, unsigned long long want, volatile unsigned long long* timeout_marker,
                            long long max_cycles) {
    const long long t0 = clock64();
    while (*flag < want) {
        // never hang the GPU: give up after max_cycles and leave a STICKY marker in host memory; later waits return at once and
        // every host entry point that touches state fails (p2p_poisoned)
        if (*timeout_marker != 0 || clock64() - t0 > max_cycles) {
            if (*timeout_marker == 0) *timeout_marker = want;
            __threadfence_system();
            return;
        }
        __nanosleep(200);
    }
    __threadfence_system();   // acquire: the planes pushed before the flag are visible to the kernels that follow
}

__global__ void k_signal_flag(volatile unsigned long long* peer_flag, unsigned long long value) {
    __threadfence_system();   // release: this stream's earlier peer writes (the pushed planes) are ordered before the flag
    *peer_flag = value;
}

float* set_base(fdtd_b200_handle h, int set, int c) { return set == h->cur_set ? h->alloc[c] : h->alt_alloc[c]; }

int p2p_open(fdtd_b200_handle h, const P2PBlob& b, fdtd_b200_solver::PeerLink* link) {
    link->present = true;
    link->nzl = b.z_end - b.z_begin;
    link->same_process = b.pid == static_cast<uint64_t>(getpid());
    if (link->same_process) {
        if (b.device != h->device) {
            int can = 0;
            CU_TRY(h, cudaDeviceCanAccessPeer(&can, h->device, b.device));
            if (!can) return fail(h, FDTD_B200_ECUDA, "devices cannot access each other's memory (no NVLink/P2P path)");
            const cudaError_t e = cudaDeviceEnablePeerAccess(b.device, 0);
            if (e != cudaSuccess && e != cudaErrorPeerAccessAlreadyEnabled) CU_TRY(h, e);
            (void)cudaGetLastError();
        }
        for (int s = 0; s < 2; ++s)
            for (int c = 0; c < 6; ++c) link->field[s][c] = reinterpret_cast<float*>(b.raw_field[s][c]);
        link->mat = reinterpret_cast<unsigned char*>(b.raw_mat);
        link->flags = reinterpret_cast<unsigned long long*>(b.raw_flags);
        link->psi_ghost = reinterpret_cast<float*>(b.raw_psi_ghost);
        return FDTD_B200_OK;
    }
    auto open = [&](const cudaIpcMemHandle_t& m, void** out) -> int {
        CU_TRY(h, cudaIpcOpenMemHandle(out, m, cudaIpcMemLazyEnablePeerAccess));
        link->opened.push_back(*out);
        return FDTD_B200_OK;
    };
    void* p = nullptr;
    for (int s = 0; s < 2; ++s)
        for (int c = 0; c < 6; ++c) {
            if (int rc = open(b.field[s][c], &p)) return rc;
            link->field[s][c] = static_cast<float*>(p);
        }
    if (int rc = open(b.mat, &p)) return rc;
    link->mat = static_cast<unsigned char*>(p);
    if (int rc = open(b.flags, &p)) return rc;
    link->flags = static_cast<unsigned long long*>(p);
    if (b.raw_psi_ghost) {
        if (int rc = open(b.psi_ghost, &p)) return rc;
        link->psi_ghost = static_cast<float*>(p);
    }
    return FDTD_B200_OK;
}

// floats per ψ_H array of this slab's first plane, per axis (what the slab below needs of it): x [y][k], y [k][x], z [y][x]
void pml_first_plane(fdtd_b200_handle h, uint64_t out[3]) {
    out[0] = static_cast<uint64_t>(h->pml[0].n) * h->ny;
    out[1] = static_cast<uint64_t>(h->pml[1].n) * h->nx;
    out[2] = (h->pml[2].n && h->pml[2].pos[0] == 0) ? h->nx * h->ny : 0;
}

void p2p_close(fdtd_b200_solver::PeerLink* link) {
    for (void* p : link->opened) cudaIpcCloseMemHandle(p);
    link->opened.clear();
    link->present = false;
}

// Store this slab's boundary planes of buffer set `set` into the neighbours' ghost planes of the same set, then raise their flags.
int p2p_push(fdtd_b200_handle h, int set, cudaStream_t st, bool stored_by_kernel) {
    const size_t plane_bytes = static_cast<size_t>(h->plane) * sizeof(float);
    if (stored_by_kernel) {   // the boundary CTAs stored the planes themselves (k_step_fused<…, PUSH>); only the one-off index plane is copied
        if (h->lower.present && !h->p2p_mat_sent)
            CU_TRY(h, cudaMemcpyAsync(h->lower.mat + h->plane * h->lower.nzl * h->idx_bytes, h->mat, h->plane * h->idx_bytes, cudaMemcpyDeviceToDevice, st));
    } else {
    if (h->upper.present) {   // my top Ex,Ey → the upper slab's ghost plane below its plane 0
        for (int c = 0; c < 2; ++c)
            CU_TRY(h, cudaMemcpyAsync(h->upper.field[set][c], set_base(h, set, c) + h->plane * h->nzl, plane_bytes, cudaMemcpyDeviceToDevice, st));
    }
    if (h->lower.present) {   // my whole first plane (E and Hx,Hy) → the lower slab's ghost plane above its last plane
        for (int c = 0; c < 5; ++c)
            CU_TRY(h, cudaMemcpyAsync(h->lower.field[set][c] + h->plane * (h->lower.nzl + 1), set_base(h, set, c) + h->plane, plane_bytes,
                                      cudaMemcpyDeviceToDevice, st));
        if (!h->p2p_mat_sent)
            CU_TRY(h, cudaMemcpyAsync(h->lower.mat + h->plane * h->lower.nzl * h->idx_bytes, h->mat, h->plane * h->idx_bytes, cudaMemcpyDeviceToDevice, st));
    }
    }
    if (h->has_pml && h->lower.present && h->lower.psi_ghost) {
        // ψ_H of my first plane → the slab below (it recomputes H' of that plane): the set that belongs to `set`'s fields, into
        // the slot of the tick those fields belong to
        const bool pending = set != h->cur_set;
        uint64_t fp[3];
        pml_first_plane(h, fp);
        const size_t slot = 2 * (fp[0] + fp[1] + fp[2]);
        float* dst = h->lower.psi_ghost + ((h->tick + (pending ? 1 : 0)) & 1) * slot;
        for (int a = 0; a < 3; ++a) {
            for (int g = 0; g < 2 && fp[a]; ++g)
                CU_TRY(h, cudaMemcpyAsync(dst + g * fp[a], pending ? h->pml[a].psi_alt[g] : h->pml[a].psi[0][g], fp[a] * sizeof(float),
                                          cudaMemcpyDeviceToDevice, st));
            dst += 2 * fp[a];
        }
    }
    h->p2p_mat_sent = true;
    h->halo_seq += 1;
    if (h->upper.present) k_signal_flag<<<1, 1, 0, st>>>(h->upper.flags + 0, h->halo_seq);   // "received from below"
    if (h->lower.present) k_signal_flag<<<1, 1, 0, st>>>(h->lower.flags + 1, h->halo_seq);   // "received from above"
    CU_TRY(h, cudaGetLastError());
    return FDTD_B200_OK;
}

int p2p_wait(fdtd_b200_handle h, cudaStream_t st) {
    if (h->halo_seq == 0) return FDTD_B200_OK;
    // ≈ 10 s at 1.9 GHz; FDTD_B200_P2P_TIMEOUT_MS shortens it (read per call: tests set it for one case)
    const char* e_ms = std::getenv("FDTD_B200_P2P_TIMEOUT_MS");
    const long long ms_env = e_ms ? std::atoll(e_ms) : 0;
    const long long max_cycles = (ms_env > 0 ? ms_env : 10000LL) * 1900000LL;
    if (h->lower.present) k_wait_flag<<<1, 1, 0, st>>>(h->d_flags + 0, h->halo_seq, h->hd_timeout, max_cycles);
    if (h->upper.present) k_wait_flag<<<1, 1, 0, st>>>(h->d_flags + 1, h->halo_seq, h->hd_timeout, max_cycles);
    CU_TRY(h, cudaGetLastError());
    return FDTD_B200_OK;
}

}  // namespace

extern "C" {

int fdtd_b200_p2p_connected(fdtd_b200_handle h) { return h && h->p2p ? 1 : 0; }

int fdtd_b200_p2p_export(fdtd_b200_handle h, void* blob) {
    if (!h || !blob) return fail(h, FDTD_B200_EINVAL, "NULL argument");
    if (!h->fused) return fail(h, FDTD_B200_ESTATE, "the peer-memory transport needs a fused-step handle; this one is not: " + h->fused_why_not);
    if (h->has_pml && !h->pml_in_kernel) return fail(h, FDTD_B200_ESTATE, "the peer-memory transport carries a PML only inside the fused kernel (fdtd_b200_pml_in_kernel); use the two-kernel slab path");
    if (int rc = set_device(h)) return rc;
    if (!h->d_flags) {
        CU_TRY(h, cudaMalloc(&h->d_flags, 64));
        CU_TRY(h, cudaMemset(h->d_flags, 0, 64));
    }
    if (!h->h_timeout) {
        void* hp = nullptr; void* dp = nullptr;
        CU_TRY(h, cudaHostAlloc(&hp, 64, cudaHostAllocMapped | cudaHostAllocPortable));
        std::memset(hp, 0, 64);
        CU_TRY(h, cudaHostGetDevicePointer(&dp, hp, 0));
        h->h_timeout = static_cast<volatile unsigned long long*>(hp);
        h->hd_timeout = static_cast<unsigned long long*>(dp);
    }
    // receive buffer for ψ_H of the plane above this slab (the x and y profiles are global, the z entry exists if that plane is a PML coordinate)
    if (h->psi_ghost) { cudaFree(h->psi_ghost); h->psi_ghost = nullptr; }
    h->ghost_plane[0] = h->ghost_plane[1] = h->ghost_plane[2] = 0;
    if (h->has_pml && h->z_end < h->nz) {
        h->ghost_plane[0] = static_cast<size_t>(h->pml[0].n) * h->ny;
        h->ghost_plane[1] = static_cast<size_t>(h->pml[1].n) * h->nx;
        h->ghost_plane[2] = h->pml[2].ghost ? h->nx * h->ny : 0;
        const size_t floats = std::max<size_t>(4 * (h->ghost_plane[0] + h->ghost_plane[1] + h->ghost_plane[2]), 1);
        CU_TRY(h, cudaMalloc(&h->psi_ghost, floats * sizeof(float)));
        CU_TRY(h, cudaMemset(h->psi_ghost, 0, floats * sizeof(float)));
    }
    P2PBlob b;
    std::memset(&b, 0, sizeof b);
    b.magic = P2P_MAGIC; b.version = 1; b.device = h->device; b.idx_bytes = static_cast<uint32_t>(h->idx_bytes);
    b.pid = static_cast<uint64_t>(getpid());
    b.nx = h->nx; b.ny = h->ny; b.nz = h->nz; b.z_begin = h->z_begin; b.z_end = h->z_end;
    for (int s = 0; s < 2; ++s)
        for (int c = 0; c < 6; ++c) {
            float* base = set_base(h, s, c);
            b.raw_field[s][c] = reinterpret_cast<uint64_t>(base);
            CU_TRY(h, cudaIpcGetMemHandle(&b.field[s][c], base));
        }
    b.raw_mat = reinterpret_cast<uint64_t>(h->mat);
    b.raw_flags = reinterpret_cast<uint64_t>(h->d_flags);
    CU_TRY(h, cudaIpcGetMemHandle(&b.mat, h->mat));
    CU_TRY(h, cudaIpcGetMemHandle(&b.flags, h->d_flags));
    if (h->psi_ghost) {
        b.raw_psi_ghost = reinterpret_cast<uint64_t>(h->psi_ghost);
        CU_TRY(h, cudaIpcGetMemHandle(&b.psi_ghost, h->psi_ghost));
    }
    for (int a = 0; a < 3; ++a) b.ghost_plane[a] = h->ghost_plane[a];
    if (h->has_pml) pml_first_plane(h, b.first_plane);
    std::memset(blob, 0, FDTD_B200_P2P_BLOB_BYTES);
    std::memcpy(blob, &b, sizeof b);
    return FDTD_B200_OK;
}

int fdtd_b200_p2p_connect(fdtd_b200_handle h, const void* lower_blob, const void* upper_blob) {
    if (!h) return fail(nullptr, FDTD_B200_EINVAL, "NULL handle");
    if (!h->fused || !h->d_flags) return fail(h, FDTD_B200_ESTATE, "call fdtd_b200_p2p_export on this (fused-step) handle first");
    if ((h->z_begin > 0) != (lower_blob != nullptr) || (h->z_end < h->nz) != (upper_blob != nullptr))
        return fail(h, FDTD_B200_EINVAL, "a neighbour blob is required exactly where the slab has a neighbour");
    if (int rc = set_device(h)) return rc;
    p2p_close(&h->lower);
    p2p_close(&h->upper);
    h->p2p = false;
    for (int side = 0; side < 2; ++side) {
        const void* src = side == 0 ? lower_blob : upper_blob;
        if (!src) continue;
        P2PBlob b;
        std::memcpy(&b, src, sizeof b);
        if (b.magic != P2P_MAGIC || b.version != 1) return fail(h, FDTD_B200_EINVAL, "not a fdtd_b200 p2p blob");
        if (b.nx != h->nx || b.ny != h->ny || b.nz != h->nz || b.idx_bytes != static_cast<uint32_t>(h->idx_bytes))
            return fail(h, FDTD_B200_EINVAL, "neighbour blob describes a different lattice");
        if (side == 0 ? b.z_end != h->z_begin : b.z_begin != h->z_end) return fail(h, FDTD_B200_EINVAL, "neighbour slab is not adjacent");
        {   // both sides must have been given the same PML: what the upper slab pushes is what the lower one expects
            uint64_t mine[3] = {0, 0, 0};
            if (h->has_pml) pml_first_plane(h, mine);
            for (int a = 0; a < 3; ++a) {
                const uint64_t expect = side == 0 ? b.ghost_plane[a] : h->ghost_plane[a];
                const uint64_t sent = side == 0 ? mine[a] : b.first_plane[a];
                if (expect != sent) return fail(h, FDTD_B200_EINVAL, "the neighbour slab was given a different PML (set the same profiles on every slab before fdtd_b200_p2p_export)");
            }
        }
        if (int rc = p2p_open(h, b, side == 0 ? &h->lower : &h->upper)) return rc;
    }
    if (!h->bstream) {
        int lo = 0, hi = 0;
        CU_TRY(h, cudaDeviceGetStreamPriorityRange(&lo, &hi));
        CU_TRY(h, cudaStreamCreateWithPriority(&h->bstream, cudaStreamNonBlocking, hi));
        CU_TRY(h, cudaEventCreateWithFlags(&h->ev_main, cudaEventDisableTiming));
        CU_TRY(h, cudaEventCreateWithFlags(&h->ev_bnd, cudaEventDisableTiming));
        CU_TRY(h, cudaEventCreateWithFlags(&h->ev_push, cudaEventDisableTiming));
    }
    h->p2p = true;
    h->p2p_mat_sent = false;
    h->halos_dirty = true;   // nothing has been pushed yet: fdtd_b200_p2p_refresh comes first
    return FDTD_B200_OK;
}

int fdtd_b200_p2p_refresh(fdtd_b200_handle h) {
    if (!h || !h->p2p) return fail(h, FDTD_B200_ESTATE, "fdtd_b200_p2p_connect has not succeeded on this handle");
    if (int rc = set_device(h)) return rc;
    if (int rc = p2p_poisoned(h)) return rc;
    // asynchronous and flag-ordered: the next p2p_step's boundary launch waits on the device for the neighbours' copies
    h->halos_dirty = false;
    return p2p_push(h, h->cur_set, h->stream, false);
}

int fdtd_b200_p2p_step(fdtd_b200_handle h, uint64_t n_steps) {
    if (!h || !h->p2p) return fail(h, FDTD_B200_ESTATE, "fdtd_b200_p2p_connect has not succeeded on this handle");
    if (int rc = set_device(h)) return rc;
    if (int rc = p2p_poisoned(h)) return rc;   // sticky: a flag wait of an earlier step timed out
    if (h->halos_dirty)
        return fail(h, FDTD_B200_ESTATE, "the neighbours' ghost planes are stale (connect, reset, set_time, write_field or write_pml_state since the last push): "
                                         "call fdtd_b200_p2p_refresh on every slab first");
    // Two slabs driven by ONE host thread: a slab's step s+1 waits on the device for its neighbour's step s, which this thread
    // has not enqueued yet — with several steps per call the launch queue fills behind that wait and the call deadlocks until
    // the flag wait times out.  One step per call and slab, slabs in turn, is the only order that cannot.
    if (n_steps > 1 && ((h->lower.present && h->lower.same_process) || (h->upper.present && h->upper.same_process)))
        return fail(h, FDTD_B200_EINVAL, "neighbour slabs in the same process: call fdtd_b200_p2p_step(h, 1) on each slab in turn");
    const int n = static_cast<int>(h->nzl);
    const HostRowScope scope(h, n_steps == 1);
    // Split of the slab into the boundary launches (high-priority stream: their stores into the neighbours' ghost planes leave
    // early in the step) and the interior launch (main stream).  The boundary launches are WHOLE CHUNKS of the march a single
    // launch over the slab would make — [0, zc) and [n - zc, n) — so the step pays exactly as many ring prologues and
    // recomputed H' planes as an unsplit step.  (Round 1 launched the two boundary PLANES on their own: a one-plane march is all
    // prologue, ≈ 3.5 plane-times each, 4 % of a 128-plane slab — the named limiter of the 1→8 weak-scaling curve.)  Slabs too
    // thin for three chunks fall back to one-plane boundary launches, slabs under three planes are all boundary.
    static const bool plane_split = [] { const char* e = std::getenv("FDTD_B200_P2P_PLANE_SPLIT"); return e && e[0] == '1'; }();
    int zc_chunk = 0;
    if (!plane_split && n >= 3) {
        constexpr int BX = 32, BYc = 7;   // tile of k_step_fused<…, 32, 8>: 128 cells × 7 core rows
        const long long tiles = static_cast<long long>((h->nx + 4 * BX - 1) / (4 * BX)) * ((h->ny + BYc - 1) / BYc);
        const int n_bnd = (h->lower.present ? 1 : 0) + (h->upper.present ? 1 : 0);
        const int zc_full = h->tune.zc > 0 ? h->tune.zc : pick_fused_zc(h, tiles, n, n_bnd + 1, h->has_pml ? 32 : 64);   // PML launches: see launch_fused_t
        if (zc_full * n_bnd < n) zc_chunk = zc_full;
    }
    const bool split = n >= 3;                      // otherwise the whole slab is "boundary"
    const int bw = zc_chunk > 0 ? zc_chunk : 1;     // planes per boundary launch
    const int lo = (split && h->lower.present) ? bw : 0;
    const int hi = (split && h->upper.present) ? n - bw : n;
    struct ZcScope { fdtd_b200_handle h; ~ZcScope() { h->zc_override = 0; } } zc_scope{h};
    h->zc_override = zc_chunk;
    cudaStream_t S = h->stream, B = h->bstream;
    // default: the boundary CTAs store into the neighbours' ghost planes themselves; FDTD_B200_P2P_COPY_ENGINE=1 selects
    // the earlier variant (seven peer cudaMemcpyAsync per step and cut after the boundary kernels) for comparison
    static const bool copy_engine = [] { const char* e = std::getenv("FDTD_B200_P2P_COPY_ENGINE"); return e && e[0] == '1'; }();
    const bool in_kernel = !copy_engine && !h->has_pml;   // the PML instantiation has no PUSH variant: its planes go by copy engine with the ψ planes
    // whatever was enqueued on the main stream before (uploads, earlier steps, a refresh) precedes the first boundary launch
    CU_TRY(h, cudaEventRecord(h->ev_main, S));
    for (uint64_t s = 0; s < n_steps; ++s) {
        // boundary stream: previous interior done → neighbours' pushes of the previous step arrived → boundary planes → push
        CU_TRY(h, cudaStreamWaitEvent(B, h->ev_main, 0));
        if (int rc = p2p_wait(h, B)) return rc;
        h->stream = B;
        int rc = FDTD_B200_OK;
        if (!split) rc = launch_fused(h, 0, n, in_kernel);
        else {
            if (hi < n) rc = launch_fused(h, hi, n, in_kernel);
            if (!rc && lo > 0) rc = launch_fused(h, 0, lo, in_kernel);
        }
        h->stream = S;
        if (rc) return rc;
        CU_TRY(h, cudaEventRecord(h->ev_bnd, B));
        if (int rc2 = p2p_push(h, h->cur_set ^ 1, B, in_kernel)) return rc2;   // the planes just written live in the pending set
        CU_TRY(h, cudaEventRecord(h->ev_push, B));
        // main stream: interior planes, concurrently with the boundary work above (they read the same current set and write
        // disjoint planes of the pending set); the previous step's boundary planes are ordered by ev_bnd from the last turn
        if (split && lo < hi) {
            if (int rc3 = launch_fused(h, lo, hi)) return rc3;
        }
        CU_TRY(h, cudaStreamWaitEvent(S, h->ev_bnd, 0));            // the next interior launch reads this step's boundary planes
        CU_TRY(h, cudaEventRecord(h->ev_main, S));
        commit_fused(h);
        end_step(h);
    }
    CU_TRY(h, cudaStreamWaitEvent(S, h->ev_push, 0));               // sync()/timer on the main stream cover the last pushes too
    if (n_steps == 1 && h->n_probes_total) return probe_mirror_row(h, h->tick - 1);   // one-step protocol, as fdtd_b200_step
    return FDTD_B200_OK;
}

}  // extern "C"

// ---- field readback without any allocation ---------------------------------------------------------------------------------------
// No counterpart in the reference: added for callers that cannot call cudaMalloc (code
// under CUDA graph capture, tight serving loops, harnesses that forbid it).  fdtd_b200_read_field gathers the box into a
// staging buffer that it (re)allocates on demand; this variant gathers into a file-scope device array of a fixed, modest size.
// Every GPU that loads the library holds the array (FDTD_B200_STATIC_SCRATCH_BYTES) until the process exits.  All calls, on all
// handles of a device, share it, and it keeps the last call's contents: the gather kernel overwrites every element it later copies
// out, so nothing has to start at zero — but two calls that overlap on the GPU (different streams / threads) race on it.
static __device__ __align__(256) float g_static_scratch[FDTD_B200_STATIC_SCRATCH_BYTES / sizeof(float)];

namespace {

// In host code `g_static_scratch` is a host-side stand-in, not a device address: the address the kernel needs comes from
// cudaGetSymbolAddress, looked up once per device with that device current.
int static_scratch_ptr(fdtd_b200_handle h, float** out) {
    constexpr int kMaxDevices = 64;
    static std::mutex mu;
    static float* ptr[kMaxDevices] = {};
    if (h->device < 0 || h->device >= kMaxDevices) return fail(h, FDTD_B200_EINVAL, "device ordinal beyond the static scratch table");
    std::lock_guard<std::mutex> lock(mu);
    if (!ptr[h->device]) {
        void* p = nullptr;
        CU_TRY(h, cudaGetSymbolAddress(&p, g_static_scratch));   // the caller made h->device current (set_device)
        if (!p) return fail(h, FDTD_B200_ECUDA, "cudaGetSymbolAddress returned a null address for the static scratch array");
        ptr[h->device] = static_cast<float*>(p);
    }
    *out = ptr[h->device];
    return FDTD_B200_OK;
}

}  // namespace

extern "C" int fdtd_b200_read_field_static(fdtd_b200_handle h, int component, const uint64_t lo[3], const uint64_t hi[3], float* out) {
    if (!h) return fail(nullptr, FDTD_B200_EINVAL, "NULL handle");
    if (component != FDTD_B200_FIELD_E && component != FDTD_B200_FIELD_H) return fail(h, FDTD_B200_EINVAL, "component must be E or H");
    BoxView B{};
    size_t n = 0;
    if (int rc = make_box(h, lo, hi, &B, &n)) return rc;
    if (n == 0) return FDTD_B200_OK;
    if (!out) return fail(h, FDTD_B200_EINVAL, "NULL output buffer");
    constexpr size_t kMaxCells = FDTD_B200_STATIC_SCRATCH_BYTES / (3 * sizeof(float));
    if (n > kMaxCells)
        return fail(h, FDTD_B200_EINVAL, "fdtd_b200_read_field_static: box of " + std::to_string(n) + " cells exceeds the static scratch array (" +
                                             std::to_string(kMaxCells) + " cells = FDTD_B200_STATIC_SCRATCH_BYTES / 12); use fdtd_b200_read_field or read the box in parts");
    if (int rc = set_device(h)) return rc;
    float* scratch = nullptr;
    if (int rc = static_scratch_ptr(h, &scratch)) return rc;
    const int c0 = component == FDTD_B200_FIELD_E ? 0 : 3;
    const int blocks = static_cast<int>(std::min<size_t>((n + 255) / 256, static_cast<size_t>(h->n_sm) * 16));
    k_gather_box<<<blocks, 256, 0, h->stream>>>(h->f[c0], h->f[c0 + 1], h->f[c0 + 2], B, scratch);
    h->launches += 1;
    CU_TRY(h, cudaGetLastError());
    CU_TRY(h, cudaMemcpyAsync(out, scratch, n * 3 * sizeof(float), cudaMemcpyDeviceToHost, h->stream));
    CU_TRY(h, cudaStreamSynchronize(h->stream));
    return p2p_poisoned(h);
}
