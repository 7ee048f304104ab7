// TEST INFRASTRUCTURE — NOT PRODUCT CODE.  C ABI over oracle/fdtd_ref.hpp so that
// pytest (ctypes) and bench.py's cpu_baseline leg can drive the CPU restatement
// of cem-solver's FDTD path.  See the header for the parity status.
#include "fdtd_ref.hpp"

#include <chrono>
#include <memory>

#ifdef _OPENMP
#include <omp.h>
#endif

using namespace fdtd_ref;

namespace {

struct Any {
    int precision;  // 0 = f64 (CPU backend), 1 = f32 (WGSL arithmetic)
    std::unique_ptr<Solver<double>> d;
    std::unique_ptr<Solver<float>> f;
};

template <class F>
auto visit(Any* a, F&& fn) {
    if (a->precision == 0) return fn(*a->d);
    return fn(*a->f);
}

}  // namespace

extern "C" {

void* fdtd_ref_create(int precision, const uint64_t n[3], const double dx[3], double dt) {
    auto* a = new Any{};
    a->precision = precision;
    const size_t size[3] = {static_cast<size_t>(n[0]), static_cast<size_t>(n[1]), static_cast<size_t>(n[2])};
    Resolution r{{dx[0], dx[1], dx[2]}, dt};
    if (precision == 0) a->d = std::make_unique<Solver<double>>(size, r);
    else a->f = std::make_unique<Solver<float>>(size, r);
    return a;
}

void fdtd_ref_destroy(void* h) { delete static_cast<Any*>(h); }

int fdtd_ref_max_threads(void) {
#ifdef _OPENMP
    return omp_get_max_threads();
#else
    return 1;
#endif
}
void fdtd_ref_set_threads(int n) {
#ifdef _OPENMP
    omp_set_num_threads(n);
#else
    (void)n;
#endif
}

// cpu/mod.rs:244-264: one UpdateCoefficients::new per linear index, in order.
// palette rows: (mu_r, sigma_m, eps_r, sigma) as material.rs:105-116 declares.
int fdtd_ref_set_materials(void* h, const double* palette, uint32_t n_palette, const uint32_t* ids, double eps0, double mu0) {
    auto* a = static_cast<Any*>(h);
    try {
        return visit(a, [&](auto& s) {
            const size_t len = s.strider.len();
            Resolution r{{0, 0, 0}, s.dt};
            PhysicalConstants pc{eps0, mu0};
            for (size_t i = 0; i < len; ++i) {
                const uint32_t id = ids ? ids[i] : 0u;
                if (id >= n_palette) return 1;
                const double* p = palette + 4 * static_cast<size_t>(id);
                s.set_coefficients(i, update_coefficients(r, pc, Material{p[0], p[1], p[2], p[3]}));
            }
            return 0;
        });
    } catch (const std::exception&) {
        return 2;
    }
}

void fdtd_ref_update_coefficients(double dt, double eps0, double mu0, const double material[4], double out[4]) {
    Resolution r{{0, 0, 0}, dt};
    const UpdateCoefficients u = update_coefficients(r, PhysicalConstants{eps0, mu0}, Material{material[0], material[1], material[2], material[3]});
    out[0] = u.c_a; out[1] = u.c_b; out[2] = u.d_a; out[3] = u.d_b;
}

// pml.rs:34-86.  out: b[3], c[3].
void fdtd_ref_pml_coefficients(double dt, double eps0, double mu0, double sigma, double kappa, double a, const double normal[3], double out[6]) {
    Resolution r{{0, 0, 0}, dt};
    const PmlCoefficients p = pml_coefficients_new(r, PhysicalConstants{eps0, mu0}, sigma, kappa, a, normal);
    for (int i = 0; i < 3; ++i) { out[i] = p.b[i]; out[3 + i] = p.c[i]; }
}
void fdtd_ref_pml_coefficients_graded(double dt, double eps0, double mu0, double m, double m_a, double sigma_max, double kappa_max, double a_max,
                                      double depth, const double normal[3], double out[6]) {
    Resolution r{{0, 0, 0}, dt};
    const PmlCoefficients p = pml_coefficients_new_graded(r, PhysicalConstants{eps0, mu0}, m, m_a, sigma_max, kappa_max, a_max, depth, normal);
    for (int i = 0; i < 3; ++i) { out[i] = p.b[i]; out[3 + i] = p.c[i]; }
}

// cpu/mod.rs:242-255: the instance asks `pml(point)` for every cell; here the caller passes the cells that answered Some.
// cells: linear indices; bc: [n][6] = b xyz, c xyz; bc_h: the same for the H half-step (NULL = the reference's model: one set).
int fdtd_ref_set_pml(void* h, uint64_t n, const uint64_t* cells, const double* bc, const double* bc_h) {
    return visit(static_cast<Any*>(h), [&](auto& s) {
        for (uint64_t k = 0; k < n; ++k) {
            if (cells[k] >= s.strider.len()) return 1;
            PmlCoefficients p;
            for (int i = 0; i < 3; ++i) { p.b[i] = bc[6 * k + i]; p.c[i] = bc[6 * k + 3 + i]; }
            PmlCoefficients ph = p;
            if (bc_h) for (int i = 0; i < 3; ++i) { ph.b[i] = bc_h[6 * k + i]; ph.c[i] = bc_h[6 * k + 3 + i]; }
            s.set_pml(static_cast<size_t>(cells[k]), p, ph);
        }
        return 0;
    });
}

void fdtd_ref_reset(void* h) { visit(static_cast<Any*>(h), [](auto& s) { s.reset_state(); return 0; }); }
void fdtd_ref_begin_update(void* h) { visit(static_cast<Any*>(h), [](auto& s) { s.begin_update(); return 0; }); }
int fdtd_ref_set_forcing(void* h, const uint64_t point[3], const double j[3], const double m[3]) {
    const size_t p[3] = {static_cast<size_t>(point[0]), static_cast<size_t>(point[1]), static_cast<size_t>(point[2])};
    SourceValues v{{j[0], j[1], j[2]}, {m[0], m[1], m[2]}};
    return visit(static_cast<Any*>(h), [&](auto& s) { return s.set_forcing(p, v) ? 0 : 1; });
}
void fdtd_ref_finish(void* h) { visit(static_cast<Any*>(h), [](auto& s) { s.finish(); return 0; }); }
uint64_t fdtd_ref_tick(void* h) { return visit(static_cast<Any*>(h), [](auto& s) { return static_cast<uint64_t>(s.tick); }); }
double fdtd_ref_time(void* h) { return visit(static_cast<Any*>(h), [](auto& s) { return s.time; }); }

// out: [cells][3] doubles (f32 state widened, as wgpu/mod.rs:522 does).
void fdtd_ref_read_field(void* h, int component, int stale, double* out) {
    visit(static_cast<Any*>(h), [&](auto& s) {
        const auto& f = s.field(component, stale != 0);
        for (size_t i = 0; i < f.size(); ++i) { out[3 * i] = f[i].x; out[3 * i + 1] = f[i].y; out[3 * i + 2] = f[i].z; }
        return 0;
    });
}

void fdtd_ref_write_field(void* h, int component, const double* in) {
    visit(static_cast<Any*>(h), [&](auto& s) {
        auto& f = s.field_mut(component);
        using S = std::decay_t<decltype(f[0].x)>;
        for (size_t i = 0; i < f.size(); ++i) { f[i].x = static_cast<S>(in[3 * i]); f[i].y = static_cast<S>(in[3 * i + 1]); f[i].z = static_cast<S>(in[3 * i + 2]); }
        return 0;
    });
}

// Batched driver loop ≙ cem-app/src/solver/runner.rs:522-526: per step
// begin_update → set_forcing(point, values(t_n)) for each source → finish, then
// sample `probe_cells` from the fresh buffers (Field::field semantics).
// jm: [n_steps][n_src][6] (j xyz, m xyz), probe_out: [n_steps][n_probes][6] (E xyz, H xyz).
// Returns wall seconds spent in the loop.
double fdtd_ref_run(void* h, uint64_t n_steps, uint32_t n_src, const uint64_t* src_points, const double* jm,
                    uint32_t n_probes, const uint64_t* probe_cells, double* probe_out) {
    const auto t0 = std::chrono::steady_clock::now();
    visit(static_cast<Any*>(h), [&](auto& s) {
        for (uint64_t step = 0; step < n_steps; ++step) {
            s.begin_update();
            for (uint32_t k = 0; k < n_src; ++k) {
                const size_t p[3] = {static_cast<size_t>(src_points[3 * k]), static_cast<size_t>(src_points[3 * k + 1]), static_cast<size_t>(src_points[3 * k + 2])};
                const double* v = jm + (step * n_src + k) * 6;
                s.set_forcing(p, SourceValues{{v[0], v[1], v[2]}, {v[3], v[4], v[5]}});
            }
            s.finish();
            if (n_probes) {
                const auto& e = s.field(0, false);
                const auto& hf = s.field(1, false);
                for (uint32_t k = 0; k < n_probes; ++k) {
                    double* o = probe_out + (step * n_probes + k) * 6;
                    const size_t c = static_cast<size_t>(probe_cells[k]);
                    o[0] = e[c].x; o[1] = e[c].y; o[2] = e[c].z;
                    o[3] = hf[c].x; o[4] = hf[c].y; o[5] = hf[c].z;
                }
            }
        }
        return 0;
    });
    return std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
}

void fdtd_ref_project(void* h, int component, int stale, const float projection[16], const float color_map[16], int mode,
                      uint32_t w, uint32_t hgt, uint8_t* rgba) {
    visit(static_cast<Any*>(h), [&](auto& s) { project_to_image(s, component, stale != 0, projection, color_map, mode, w, hgt, rgba); return 0; });
}

// util.rs:130-209; out receives up to cap points (3 u64 each); returns the count.
uint64_t fdtd_ref_iter_points(const uint64_t start[3], const uint64_t end[3], uint64_t* out, uint64_t cap) {
    const size_t s[3] = {static_cast<size_t>(start[0]), static_cast<size_t>(start[1]), static_cast<size_t>(start[2])};
    const size_t e[3] = {static_cast<size_t>(end[0]), static_cast<size_t>(end[1]), static_cast<size_t>(end[2])};
    const auto pts = iter_points(s, e);
    const uint64_t n = pts.size() / 3;
    for (uint64_t i = 0; i < n && i < cap; ++i) { out[3 * i] = pts[3 * i]; out[3 * i + 1] = pts[3 * i + 1]; out[3 * i + 2] = pts[3 * i + 2]; }
    return n;
}

// strider.rs:34-53
void fdtd_ref_strider_point(const uint64_t n[3], uint64_t index, uint64_t out[3]) {
    const size_t size[3] = {static_cast<size_t>(n[0]), static_cast<size_t>(n[1]), static_cast<size_t>(n[2])};
    Strider st(size);
    size_t p[3];
    st.point_unchecked(static_cast<size_t>(index), p);
    out[0] = p[0]; out[1] = p[1]; out[2] = p[2];
}

// SURVEY §8 a4.  per_cell: [n][4] doubles.  Returns palette size (palette_out gets up to cap rows).
uint32_t fdtd_ref_index_materials(const double* per_cell, uint64_t n, uint32_t* ids_out, double* palette_out, uint32_t cap) {
    const auto pal = index_materials(reinterpret_cast<const Material*>(per_cell), static_cast<size_t>(n), ids_out);
    for (uint32_t k = 0; k < pal.size() && k < cap; ++k) {
        palette_out[4 * k] = pal[k].relative_permeability;
        palette_out[4 * k + 1] = pal[k].magnetic_conductivity;
        palette_out[4 * k + 2] = pal[k].relative_permittivity;
        palette_out[4 * k + 3] = pal[k].eletrical_conductivity;
    }
    return static_cast<uint32_t>(pal.size());
}

double fdtd_ref_gaussian_pulse(double time, double t0, double duration) { return gaussian_pulse(time, t0, duration); }
double fdtd_ref_continous_wave(double time, double phase, double frequency) { return continous_wave(time, phase, frequency); }

}  // extern "C"
