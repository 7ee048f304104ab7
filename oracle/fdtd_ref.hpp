// TEST INFRASTRUCTURE — NOT PRODUCT CODE.
//
// CPU restatement ("oracle") of the FDTD time-stepping path of jgraef/fdtd's
// `cem-solver` crate.  Only tests/, __graft_entry__.smoke() and bench.py's
// cpu_baseline / --impl reference legs may use anything in oracle/.
//
// PARITY STATUS: the reference is Rust; neither cargo nor rustc exist in this
// environment and the reference ships no golden vectors for the FDTD update
// (its only two tests in this crate pin point-iteration order, restated in
// tests/test_oracle.py).  So this oracle is pinned by (a) those two tests,
// (b) line-by-line correspondence with the cited reference source, and
// (c) analytic checks (decay rate, degenerate lattices).  For the field
// update itself: "parity unpinned" against reference *outputs*.
//
// Two instantiations:
//   T = double  ≙ FdtdCpuBackend        (cem-solver/src/fdtd/cpu/mod.rs)
//   T = float   ≙ the WGSL compute path (cem-solver/src/fdtd/wgpu/update.wgsl)
// Data layout deliberately mirrors the reference CPU backend: AoS Vec3<T>,
// double-buffered E/H, one 4-scalar coefficient record per cell, a per-cell
// `usize` source index and a source buffer whose slot 0 is a dummy.
#pragma once

#include <algorithm>
#include <cmath>
#include <cstddef>
#include <cstdint>
#include <cstring>
#include <stdexcept>
#include <unordered_map>
#include <vector>

namespace fdtd_ref {

template <class T>
struct Vec3 {
    T x{}, y{}, z{};
    T& operator[](int i) { return i == 0 ? x : (i == 1 ? y : z); }
    const T& operator[](int i) const { return i == 0 ? x : (i == 1 ? y : z); }
};
template <class T> inline Vec3<T> operator-(const Vec3<T>& a, const Vec3<T>& b) { return {a.x - b.x, a.y - b.y, a.z - b.z}; }
template <class T> inline Vec3<T> operator+(const Vec3<T>& a, const Vec3<T>& b) { return {a.x + b.x, a.y + b.y, a.z + b.z}; }
template <class T> inline Vec3<T> operator-(const Vec3<T>& a) { return {-a.x, -a.y, -a.z}; }
template <class T> inline Vec3<T> operator*(T s, const Vec3<T>& a) { return {s * a.x, s * a.y, s * a.z}; }
template <class T> inline Vec3<T> operator/(const Vec3<T>& a, T s) { return {a.x / s, a.y / s, a.z / s}; }

// cem-solver/src/material.rs:21-55
struct PhysicalConstants {
    double vacuum_permittivity;
    double vacuum_permeability;
    double speed_of_light() const { return std::pow(vacuum_permittivity * vacuum_permeability, -0.5); }
};

// cem-solver/src/material.rs:105-123 (field order as declared there)
struct Material {
    double relative_permeability;
    double magnetic_conductivity;
    double relative_permittivity;
    double eletrical_conductivity;
};

// cem-solver/src/fdtd/mod.rs:64-67
struct Resolution {
    double spatial[3];
    double temporal;
};

// cem-solver/src/fdtd/util.rs:91-128 — always derived in f64.
struct UpdateCoefficients {
    double c_a, c_b, d_a, d_b;
};

inline void c_or_d(double perm, double sigma, double dt, double& a, double& b) {
    const double half_sigma_dt_over_perm = 0.5 * sigma * dt / perm;
    a = (1.0 - half_sigma_dt_over_perm) / (1.0 + half_sigma_dt_over_perm);
    b = dt / (perm * (1.0 + half_sigma_dt_over_perm));
    if (std::isnan(a) || std::isnan(b)) throw std::runtime_error("UpdateCoefficients: NaN (util.rs:112-113 asserts)");
}

inline UpdateCoefficients update_coefficients(const Resolution& r, const PhysicalConstants& pc, const Material& m) {
    UpdateCoefficients u{};
    c_or_d(m.relative_permittivity * pc.vacuum_permittivity, m.eletrical_conductivity, r.temporal, u.c_a, u.c_b);
    c_or_d(m.relative_permeability * pc.vacuum_permeability, m.magnetic_conductivity, r.temporal, u.d_a, u.d_b);
    return u;
}

// cem-solver/src/fdtd/strider.rs:19-66, 147-154 — x fastest.
struct Strider {
    size_t size[3];
    size_t strides[4];
    explicit Strider(const size_t n[3]) {
        size[0] = n[0]; size[1] = n[1]; size[2] = n[2];
        strides[0] = 1;
        strides[1] = strides[0] * n[0];
        strides[2] = strides[1] * n[1];
        strides[3] = strides[2] * n[2];
    }
    size_t len() const { return strides[3]; }
    void point_unchecked(size_t index, size_t p[3]) const {  // strider.rs:34-41
        p[2] = index / strides[2];
        index %= strides[2];
        p[1] = index / strides[1];
        index %= strides[1];
        p[0] = index / strides[0];
    }
    bool is_inside(const size_t p[3]) const { return p[0] < size[0] && p[1] < size[1] && p[2] < size[2]; }
    size_t index_unchecked(const size_t p[3]) const { return p[0] * strides[0] + p[1] * strides[1] + p[2] * strides[2]; }
};

// cem-solver/src/fdtd/util.rs:130-209 — point iteration over a half-open box.
// `start`/`end` are the already-normalised bounds (normalize_point_bounds,
// util.rs:140-162: end = max(start, end) per axis).  Reproduces PointIter::next
// literally, including that the first point is emitted whenever start != end.
inline std::vector<size_t> iter_points(const size_t start_in[3], const size_t end_in[3]) {
    size_t x0[3], x1[3];
    for (int i = 0; i < 3; ++i) { x0[i] = start_in[i]; x1[i] = std::max(start_in[i], end_in[i]); }
    std::vector<size_t> out;
    bool have = !(x0[0] == x1[0] && x0[1] == x1[1] && x0[2] == x1[2]);
    size_t x[3] = {x0[0], x0[1], x0[2]};
    while (have) {
        out.push_back(x[0]); out.push_back(x[1]); out.push_back(x[2]);
        x[0] += 1;
        if (x[0] >= x1[0]) {
            x[0] = x0[0];
            x[1] += 1;
            if (x[1] >= x1[1]) {
                x[1] = x0[1];
                x[2] += 1;
                if (x[2] >= x1[2]) have = false;
            }
        }
    }
    return out;
}

// cem-solver/src/source.rs:10-13
struct SourceValues {
    double j[3];
    double m[3];
};

// cem-solver/src/source.rs:39-56
inline double gaussian_pulse(double time, double t0, double duration) {
    const double q = (time - t0) / duration;
    return std::exp(-(q * q));  // powi(2)
}
// cem-solver/src/source.rs:59-76
inline double continous_wave(double time, double phase, double frequency) {
    constexpr double TAU = 6.283185307179586476925286766559;
    return std::cos(TAU * frequency * time + phase);
}

// cem-solver/src/fdtd/pml.rs:27-86.  `PmlCoefficients { b, c }`: the scalar b (CE 7.102) and c (CE 7.99) times the
// `normal` vector the caller passes (the application passes −GradedPml.normal, runner.rs:667 — signed).
struct PmlCoefficients {
    double b[3] = {0, 0, 0};
    double c[3] = {0, 0, 0};
};

inline PmlCoefficients pml_coefficients_new(const Resolution& r, const PhysicalConstants& pc, double sigma, double kappa, double a,
                                            const double normal[3]) {
    const double b = std::exp(-((sigma / (pc.vacuum_permittivity * kappa) + a / pc.vacuum_permittivity) * r.temporal));
    const double c = sigma * (b - 1.0) / (sigma * kappa + (kappa * kappa) * a);  // kappa.powi(2)
    PmlCoefficients out;
    for (int i = 0; i < 3; ++i) { out.b[i] = b * normal[i]; out.c[i] = c * normal[i]; }
    return out;
}

// pml.rs:58-86: g1 = depth^m, g2 = (1−depth)^m_a, σ = σmax·g1, κ = 1 + (κmax−1)·g1, a = amax·g2.
inline PmlCoefficients pml_coefficients_new_graded(const Resolution& r, const PhysicalConstants& pc, double m, double m_a, double sigma_max,
                                                   double kappa_max, double a_max, double depth, const double normal[3]) {
    const double g1 = std::pow(depth, m);
    const double g2 = std::pow(1.0 - depth, m_a);
    return pml_coefficients_new(r, pc, sigma_max * g1, 1.0 + (kappa_max - 1.0) * g1, a_max * g2, normal);
}

template <class T>
class Solver {
public:
    // Per-cell coefficient record.  cpu/mod.rs:228 keeps 4×f64; the WGSL path
    // keeps vec4f cast once from f64 (wgpu/mod.rs:670-678).
    struct Coeff { T c_a, c_b, d_a, d_b; };

    Strider strider;
    T resolution[3];          // update.wgsl:159,171 divide by config.resolution[axis] (f32); cpu/util.rs:193 by f64
    double dt;
    std::vector<Coeff> coeffs;
    std::vector<Vec3<T>> e_field[2], h_field[2];  // SwapBuffer, util.rs:27-68
    std::vector<size_t> source_field;             // cpu/mod.rs:302
    struct SourceSlot { size_t cell; Vec3<T> j, m; };
    std::vector<SourceSlot> source_buffer;        // slot 0 = dummy (cpu/mod.rs:380-381)
    size_t tick = 0;
    double time = 0.0;

    // CPML.  The reference builds this storage (cpu/mod.rs:242-255 `PmlInstance{indirection, coefficients}`,
    // :649-679 `PmlState{psi_e, psi_h}`, entry 0 = "all non-pml cells") but never runs the recursion (`psi = zeros`,
    // cpu/mod.rs:419-425, 454, 498).  Here it is run, parity unpinned: per PML cell and derivative axis a the auxiliary vector
    //   G_a ← b_a·G_a + c_a·∂F/∂x_a            (CE 7.98/7.105 recursive convolution, coefficients of pml.rs),
    // and the term the reference's update already has a slot for is psi = ±curl(G_x, G_y, G_z) — the matrix form of the
    // commented-out `psi_total` (pml.rs:89-92): + for the E half-step, − for the H half-step (which subtracts the curl).
    struct PmlCell { T b[3], c[3], bh[3], ch[3]; };  // E half-step uses (b, c), H half-step (bh, ch); the reference has one set for both
    struct PsiCell { Vec3<T> g[3]; };
    std::vector<uint32_t> pml_indirection;   // per lattice cell, 0 = not a PML cell
    std::vector<PmlCell> pml_cells;          // entry 0 = dummy
    std::vector<PsiCell> psi_e, psi_h;

    void set_pml(size_t index, const PmlCoefficients& p) { set_pml(index, p, p); }
    void set_pml(size_t index, const PmlCoefficients& p, const PmlCoefficients& ph) {
        if (pml_indirection.empty()) {
            pml_indirection.assign(strider.len(), 0u);
            pml_cells.assign(1, PmlCell{});
        }
        PmlCell cell;
        for (int i = 0; i < 3; ++i) {
            cell.b[i] = static_cast<T>(p.b[i]); cell.c[i] = static_cast<T>(p.c[i]);
            cell.bh[i] = static_cast<T>(ph.b[i]); cell.ch[i] = static_cast<T>(ph.c[i]);
        }
        if (pml_indirection[index] == 0) {
            pml_indirection[index] = static_cast<uint32_t>(pml_cells.size());
            pml_cells.push_back(cell);
        } else {
            pml_cells[pml_indirection[index]] = cell;
        }
        psi_e.resize(pml_cells.size());
        psi_h.resize(pml_cells.size());
    }

    // returns curl(G_x, G_y, G_z) after advancing the recursion of PML cell `k`
    Vec3<T> advance_psi(std::vector<PsiCell>& psi, bool h_step, uint32_t k, const Vec3<T>& ddx, const Vec3<T>& ddy, const Vec3<T>& ddz) {
        const PmlCell& p = pml_cells[k];
        const T* b = h_step ? p.bh : p.b;
        const T* c = h_step ? p.ch : p.c;
        PsiCell& s = psi[k];
        s.g[0] = b[0] * s.g[0] + c[0] * ddx;
        s.g[1] = b[1] * s.g[1] + c[1] * ddy;
        s.g[2] = b[2] * s.g[2] + c[2] * ddz;
        return curl(s.g[0], s.g[1], s.g[2]);
    }

    Solver(const size_t n[3], const Resolution& res) : strider(n), dt(res.temporal) {
        for (int i = 0; i < 3; ++i) resolution[i] = static_cast<T>(res.spatial[i]);
        const size_t len = strider.len();
        coeffs.resize(len);
        for (int b = 0; b < 2; ++b) { e_field[b].assign(len, Vec3<T>{}); h_field[b].assign(len, Vec3<T>{}); }
        source_field.assign(len, 0);
    }

    void set_coefficients(size_t index, const UpdateCoefficients& u) {
        coeffs[index] = Coeff{static_cast<T>(u.c_a), static_cast<T>(u.c_b), static_cast<T>(u.d_a), static_cast<T>(u.d_b)};
    }

    // cpu/mod.rs:310-322 — fresh zeroed state.
    void reset_state() {
        for (int b = 0; b < 2; ++b) {
            std::fill(e_field[b].begin(), e_field[b].end(), Vec3<T>{});
            std::fill(h_field[b].begin(), h_field[b].end(), Vec3<T>{});
        }
        std::fill(source_field.begin(), source_field.end(), size_t(0));
        source_buffer.clear();
        std::fill(psi_e.begin(), psi_e.end(), PsiCell{});
        std::fill(psi_h.begin(), psi_h.end(), PsiCell{});
        tick = 0;
        time = 0.0;
    }

    // cpu/mod.rs:370-384
    void begin_update() {
        for (const auto& s : source_buffer) source_field[s.cell] = 0;
        source_buffer.clear();
        source_buffer.push_back(SourceSlot{0, {}, {}});
    }

    // cpu/mod.rs:391-409 — returns false where the reference panics.
    bool set_forcing(const size_t point[3], const SourceValues& v) {
        if (!strider.is_inside(point)) return false;
        const size_t cell = strider.index_unchecked(point);
        SourceSlot slot{cell,
                        {static_cast<T>(v.j[0]), static_cast<T>(v.j[1]), static_cast<T>(v.j[2])},
                        {static_cast<T>(v.m[0]), static_cast<T>(v.m[1]), static_cast<T>(v.m[2])}};
        size_t& si = source_field[cell];
        if (si == 0) {
            si = source_buffer.size();
            source_buffer.push_back(slot);
        } else {
            source_buffer[si] = slot;  // last write wins
        }
        return true;
    }

    // cpu/util.rs:168-194 with NeumannBoundaryCondition{alpha=0,beta=0}
    // (boundary_condition.rs:29-37, 70-80).  Returns the whole ∂F/∂x_axis vector.
    Vec3<T> partial_derivative(int axis, const size_t x[3], size_t dx0, size_t dx1,
                               const std::vector<Vec3<T>>& lattice) const {
        const T dx = resolution[axis];
        bool has0 = x[axis] >= dx0;
        Vec3<T> f0{}, f1{};
        if (has0) {
            size_t p[3] = {x[0], x[1], x[2]};
            p[axis] -= dx0;
            has0 = strider.is_inside(p);
            if (has0) f0 = lattice[strider.index_unchecked(p)];
        }
        size_t q[3] = {x[0], x[1], x[2]};
        q[axis] += dx1;
        const bool has1 = strider.is_inside(q);
        if (has1) f1 = lattice[strider.index_unchecked(q)];
        Vec3<T> df{};
        if (has0 && has1) df = f1 - f0;
        // (None, _) → alpha = 0 ; (_, None) → beta = 0
        return df / dx;
    }

    // cpu/util.rs:126-132: matrix columns are ∂F/∂x, ∂F/∂y, ∂F/∂z; m_rc = (∂F/∂x_c)[r]
    static Vec3<T> curl(const Vec3<T>& ddx, const Vec3<T>& ddy, const Vec3<T>& ddz) {
        return {ddy.z - ddz.y, ddz.x - ddx.z, ddx.y - ddy.x};
    }

    // cpu/mod.rs:416-515 (== update.wgsl:56-129 for T=float).
    void finish() {
        const int previous = static_cast<int>(tick % 2);  // SwapBufferIndex::from_tick
        const int next = (previous + 1) % 2;
        const long long len = static_cast<long long>(strider.len());
        const bool has_pml = !pml_indirection.empty();  // without PML cells psi is zero, as the reference executes it

        {
            const auto& e_prev = e_field[previous];
            const auto& h_prev = h_field[previous];
            auto& h_next = h_field[next];
#pragma omp parallel for schedule(static)
            for (long long i = 0; i < len; ++i) {
                size_t p[3];
                strider.point_unchecked(static_cast<size_t>(i), p);
                const Vec3<T> ddx = partial_derivative(0, p, 1, 0, e_prev);
                const Vec3<T> ddy = partial_derivative(1, p, 1, 0, e_prev);
                const Vec3<T> ddz = partial_derivative(2, p, 1, 0, e_prev);
                const Vec3<T> e_curl = curl(ddx, ddy, ddz);
                const size_t sid = source_field[i];
                Vec3<T> m_source{};
                if (sid != 0) m_source = source_buffer[sid].m;
                const Coeff& c = coeffs[i];
                Vec3<T> psi{};
                if (has_pml && pml_indirection[i] != 0) psi = -advance_psi(psi_h, true, pml_indirection[i], ddx, ddy, ddz);
                h_next[i] = c.d_a * h_prev[i] + c.d_b * (-e_curl - m_source + psi);
            }
        }
        {
            const auto& e_prev = e_field[previous];
            const auto& h_new = h_field[next];  // the H just written (cpu/mod.rs:482-484)
            auto& e_next = e_field[next];
#pragma omp parallel for schedule(static)
            for (long long i = 0; i < len; ++i) {
                size_t p[3];
                strider.point_unchecked(static_cast<size_t>(i), p);
                const Vec3<T> ddx = partial_derivative(0, p, 0, 1, h_new);
                const Vec3<T> ddy = partial_derivative(1, p, 0, 1, h_new);
                const Vec3<T> ddz = partial_derivative(2, p, 0, 1, h_new);
                const Vec3<T> h_curl = curl(ddx, ddy, ddz);
                const size_t sid = source_field[i];
                Vec3<T> j_source{};
                if (sid != 0) j_source = source_buffer[sid].j;
                const Coeff& c = coeffs[i];
                Vec3<T> psi{};
                if (has_pml && pml_indirection[i] != 0) psi = advance_psi(psi_e, false, pml_indirection[i], ddx, ddy, ddz);
                e_next[i] = c.c_a * e_prev[i] + c.c_b * (h_curl - j_source + psi);
            }
        }
        tick += 1;
        time += dt;  // accumulated, cpu/mod.rs:513
    }

    // Field::field reads buffer tick%2 (cpu/mod.rs:538); the projection pass reads
    // (tick+1)%2 (cpu/project.rs:107) — `stale` selects the latter.
    const std::vector<Vec3<T>>& field(int component /*0=E,1=H*/, bool stale = false) const {
        const int b = static_cast<int>((tick + (stale ? 1 : 0)) % 2);
        return component == 0 ? e_field[b] : h_field[b];
    }
    std::vector<Vec3<T>>& field_mut(int component) {
        const int b = static_cast<int>(tick % 2);
        return component == 0 ? e_field[b] : h_field[b];
    }
};

// sRGB OETF as used by palette 0.7 `LinSrgba → Srgba` (cpu/project.rs:154-157).
inline float linear_to_srgb(float c) {
    return c <= 0.0031308f ? 12.92f * c : 1.055f * std::pow(c, 1.0f / 2.4f) - 0.055f;
}
inline uint8_t unit_to_u8(float c) {
    float v = std::round(c * 255.0f);
    v = std::min(255.0f, std::max(0.0f, v));
    return static_cast<uint8_t>(v);
}

// cem-solver/src/fdtd/cpu/project.rs:117-165.  `projection` and `color_map` are
// column-major 4×4 (nalgebra storage).  mode 0: CPU rule (uv=(x,y)/(w+1,h+1), y
// flipped, sRGB-encoded).  mode 1: WGSL rule (project.wgsl:46-75: pixel-centre uv,
// y up, clamp(M·v,0,1) written unencoded).  mode 2: the app's `value.z` red/green
// snippet (cem-app/src/solver/runner.rs:761-780), pixel-centre uv.
template <class T>
void project_to_image(const Solver<T>& s, int component, bool stale, const float projection[16],
                      const float color_map[16], int mode, uint32_t w, uint32_t h, uint8_t* rgba) {
    const auto& field = s.field(component, stale);
    for (uint32_t y = 0; y < h; ++y) {
        for (uint32_t x = 0; x < w; ++x) {
            float u, v;
            if (mode == 0) {
                u = static_cast<float>(x) / static_cast<float>(w + 1);
                v = 1.0f - static_cast<float>(y) / static_cast<float>(h + 1);
            } else {
                u = (static_cast<float>(x) + 0.5f) / static_cast<float>(w);
                v = 1.0f - (static_cast<float>(y) + 0.5f) / static_cast<float>(h);
            }
            size_t p[3];
            for (int r = 0; r < 3; ++r) {
                // (M · (u, v, 0, 1))[r]
                const float c = projection[0 * 4 + r] * u + projection[1 * 4 + r] * v + projection[2 * 4 + r] * 0.0f + projection[3 * 4 + r];
                const float n = static_cast<float>(s.strider.size[r]);
                // cpu/project.rs:139-141: round, max(0) (NaN → 0 through `as usize`), min(s-1).
                // The WGSL path converts with vec3u(round(..)) (negative → 0) and has no upper
                // clamp (out-of-bounds read); all modes clamp here — see DESIGN.md.
                float t = std::round(c * (n - 1.0f));
                if (!(t > 0.0f)) t = 0.0f;
                p[r] = t >= n - 1.0f ? s.strider.size[r] - 1 : static_cast<size_t>(t);
            }
            uint8_t* px = rgba + 4 * (static_cast<size_t>(y) * w + x);
            const Vec3<T>& f = field[s.strider.index_unchecked(p)];
            const float fv[4] = {static_cast<float>(f.x), static_cast<float>(f.y), static_cast<float>(f.z), 1.0f};
            float col[4];
            if (mode == 2) {
                const float s_c = 10.0f, s_a = 100.0f;
                col[0] = col[1] = col[2] = col[3] = 0.0f;
                const float xz = fv[2];
                if (xz > 0.0f) { col[0] = std::min(s_c * xz, 1.0f); col[3] = std::min(s_a * xz, 1.0f); }
                else { col[1] = std::min(-s_c * xz, 1.0f); col[3] = std::min(-s_a * xz, 1.0f); }
            } else {
                for (int r = 0; r < 4; ++r) {
                    float acc = 0.0f;
                    for (int c = 0; c < 4; ++c) acc += color_map[c * 4 + r] * fv[c];
                    col[r] = std::min(1.0f, std::max(0.0f, acc));
                }
            }
            if (mode == 0) {
                // palette: colour channels get the OETF, alpha stays linear
                px[0] = unit_to_u8(linear_to_srgb(col[0]));
                px[1] = unit_to_u8(linear_to_srgb(col[1]));
                px[2] = unit_to_u8(linear_to_srgb(col[2]));
                px[3] = unit_to_u8(col[3]);
            } else {
                for (int r = 0; r < 4; ++r) px[r] = unit_to_u8(col[r]);
            }
        }
    }
}

// SURVEY §8 a4 (README.md:25-28 TODO): canonical material index — ids handed out
// in first-occurrence order over the linear index, key = the bit pattern of the
// four f64 material fields.  Returns the palette; writes ids.
inline std::vector<Material> index_materials(const Material* per_cell, size_t n, uint32_t* ids_out) {
    struct Key { uint64_t w[4]; bool operator==(const Key& o) const { return std::memcmp(w, o.w, sizeof w) == 0; } };
    struct Hash { size_t operator()(const Key& k) const { uint64_t h = 1469598103934665603ull; for (uint64_t v : k.w) { h ^= v; h *= 1099511628211ull; } return static_cast<size_t>(h); } };
    std::unordered_map<Key, uint32_t, Hash> seen;
    std::vector<Material> palette;
    for (size_t i = 0; i < n; ++i) {
        Key k;
        std::memcpy(k.w, &per_cell[i], sizeof k.w);
        auto it = seen.find(k);
        if (it == seen.end()) {
            it = seen.emplace(k, static_cast<uint32_t>(palette.size())).first;
            palette.push_back(per_cell[i]);
        }
        ids_out[i] = it->second;
    }
    return palette;
}

}  // namespace fdtd_ref
