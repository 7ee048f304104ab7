"""The Rust shim crate (rust/cem-solver-b200) cannot be compiled in this image (no cargo/rustc), so its FFI layer is checked
textually against include/fdtd_b200.h: every `pub fn` of the `extern "C"` block must exist in the header with the same
arity and argument/return types, every header function must be declared, every `#[repr(C)]` struct must list the same
fields in the same order with the same size, and every constant must carry the header's value.  The trait layer is checked
for completeness: each trait of cem-solver/src/lib.rs:28-142 and cem-solver/src/project.rs:75-133 that the runner needs
(cem-app/src/solver/runner.rs:247-259) has an `impl` in the tree, and nothing it uses is left undefined."""
import os
import re

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
CRATE = os.path.join(ROOT, "rust", "cem-solver-b200")


def _strip_c_comments(txt: str) -> str:
    return re.sub(r"/\*.*?\*/", "", txt, flags=re.S)


def _strip_rust_comments(txt: str) -> str:
    return re.sub(r"//[^\n]*", "", txt)


# C type -> the Rust spelling src/ffi.rs must use
C_TO_RUST = {
    "int": "c_int", "void": "()", "double": "f64", "float": "f32", "size_t": "usize", "char": "c_char",
    "uint64_t": "u64", "uint32_t": "u32", "int32_t": "i32", "uint8_t": "u8",
    "fdtd_b200_handle": "fdtd_b200_handle",
}
SIZES = {"f64": 8, "f32": 4, "u64": 8, "u32": 4, "i32": 4, "u8": 1}


def c_type_to_rust(ctype: str, is_array: bool) -> str:
    """'const fdtd_b200_config*' -> '*const fdtd_b200_config'; array parameters decay to pointers."""
    t = ctype.strip()
    const = t.startswith("const ")
    if const:
        t = t[len("const "):].strip()
    stars = t.count("*")
    base = t.replace("*", "").strip()
    rust = C_TO_RUST.get(base, base)   # struct names are spelled the same on both sides
    if is_array:
        stars += 1
    if stars == 0:
        return rust
    if base == "void":
        rust = "c_void"
    out = rust
    for level in range(stars):
        innermost = level == 0
        out = ("*const " if (const and innermost) else "*mut ") + out
    return out


def parse_header():
    txt = _strip_c_comments(open(os.path.join(ROOT, "include", "fdtd_b200.h")).read())
    funcs = {}
    for ret, name, args in re.findall(r"\n([a-zA-Z_][\w \*]*?)\b(fdtd_b200_\w+)\s*\(([^;{]*?)\)\s*;", txt):
        params = []
        args = " ".join(args.split())
        if args and args != "void":
            for a in args.split(","):
                a = a.strip()
                m = re.match(r"^(.*?)(\w+)(\[\d*\])?$", a)
                params.append((m.group(2), c_type_to_rust(m.group(1), bool(m.group(3)))))
        funcs[name] = (c_type_to_rust(ret, False), params)
    structs = {}
    for body, name in re.findall(r"typedef struct \w+ \{(.*?)\}\s*(\w+);", txt, flags=re.S):
        fields = []
        for decl in body.split(";"):
            decl = " ".join(decl.split())
            if not decl:
                continue
            ctype, rest = decl.split(" ", 1)
            for item in rest.split(","):
                item = item.strip()
                m = re.match(r"^(\w+)(?:\[(\d+)\])?$", item)
                rt = C_TO_RUST[ctype]
                fields.append((m.group(1), f"[{rt}; {m.group(2)}]" if m.group(2) else rt))
        structs[name] = fields
    consts = {}
    for name, val in re.findall(r"#define (FDTD_B200_\w+) (\d+)", txt):
        consts[name] = int(val)
    for body in re.findall(r"enum \{(.*?)\};", txt, flags=re.S):
        nxt = 0
        for item in body.split(","):
            item = " ".join(item.split())
            if not item:
                continue
            if "=" in item:
                k, v = [s.strip() for s in item.split("=")]
                v = v.rstrip("u")
                m = re.match(r"^(\d+)u? << (\d+)$", v)
                val = (int(m.group(1)) << int(m.group(2))) if m else int(v, 0)
            else:
                k, val = item, nxt
            consts[k] = val
            nxt = val + 1
    return funcs, structs, consts


def parse_ffi_rs():
    txt = _strip_rust_comments(open(os.path.join(CRATE, "src", "ffi.rs")).read())
    block = re.search(r'unsafe extern "C" \{(.*)\}', txt, flags=re.S).group(1)
    funcs = {}
    for name, args, ret in re.findall(r"pub fn (\w+)\s*\((.*?)\)\s*(?:->\s*([^;]+?))?\s*;", block, flags=re.S):
        params = []
        for a in [p.strip() for p in " ".join(args.split()).split(",") if p.strip()]:
            pname, ptype = a.split(":", 1)
            params.append((pname.strip().replace("r#", ""), ptype.strip()))
        funcs[name] = ((ret or "()").strip(), params)
    structs = {}
    for attrs, name, body in re.findall(r"((?:#\[[^\]]*\]\s*)+)pub struct (\w+) \{(.*?)\}", txt, flags=re.S):
        assert "repr(C)" in attrs, f"struct {name} crosses the ABI without #[repr(C)]"
        fields = []
        for f in body.split(","):
            f = " ".join(f.split())
            if f:
                fname, ftype = f.replace("pub ", "").split(":", 1)
                fields.append((fname.strip(), ftype.strip()))
        structs[name] = fields
    consts = {}
    for name, val in re.findall(r"pub const (FDTD_B200_\w+): \w+ = ([^;]+);", txt):
        m = re.match(r"^(\d+) << (\d+)$", val.strip())
        consts[name] = (int(m.group(1)) << int(m.group(2))) if m else int(val.strip(), 0)
    return funcs, structs, consts


def rust_size(t: str) -> int:
    m = re.match(r"^\[(\w+); (\d+)\]$", t)
    return SIZES[m.group(1)] * int(m.group(2)) if m else SIZES[t]


def test_every_header_function_is_declared_with_the_same_signature():
    hf, _, _ = parse_header()
    rf, _, _ = parse_ffi_rs()
    assert len(hf) >= 60
    assert set(rf) == set(hf), f"only in ffi.rs: {sorted(set(rf) - set(hf))}; only in the header: {sorted(set(hf) - set(rf))}"
    for name, (ret, params) in hf.items():
        r_ret, r_params = rf[name]
        assert r_ret == ret, f"{name}: returns {r_ret} in ffi.rs, {ret} in the header"
        assert len(r_params) == len(params), f"{name}: arity {len(r_params)} vs {len(params)}"
        for (hn, ht), (rn, rt) in zip(params, r_params):
            assert rt == ht, f"{name}({hn}): {rt} in ffi.rs, header says {ht}"
            assert rn == hn, f"{name}: parameter named {rn} in ffi.rs, {hn} in the header"


def test_repr_c_structs_match_field_for_field(b200):
    import ctypes as C

    from fdtd_b200 import _lib as L

    _, hs, _ = parse_header()
    _, rs, _ = parse_ffi_rs()
    rs = {k: v for k, v in rs.items() if k != "fdtd_b200_solver"}   # the opaque handle target
    assert set(rs) == set(hs), (sorted(rs), sorted(hs))
    ctypes_of = {"fdtd_b200_material": L.Material, "fdtd_b200_physical_constants": L.PhysicalConstants,
                 "fdtd_b200_update_coefficients": L.UpdateCoefficients, "fdtd_b200_coefficients_f32": L.CoefficientsF32,
                 "fdtd_b200_config": L.Config, "fdtd_b200_source": L.Source, "fdtd_b200_graded_pml": L.GradedPml,
                 "fdtd_b200_shape": L.Shape}
    for name, fields in hs.items():
        assert rs[name] == fields, f"{name}: {rs[name]} vs header {fields}"
        # natural alignment without padding holes: the Rust size is the sum of the fields and equals the ctypes mirror's
        size = sum(rust_size(t) for _, t in fields)
        assert size == C.sizeof(ctypes_of[name]), f"{name}: {size} bytes vs {C.sizeof(ctypes_of[name])} in the ctypes binding"
        assert [f for f, _ in fields] == [f for f, _ in ctypes_of[name]._fields_]


def test_constants_carry_the_header_values():
    _, _, hc = parse_header()
    _, _, rc = parse_ffi_rs()
    for name, val in rc.items():
        assert name in hc, f"{name} is not in the header"
        assert hc[name] == val, f"{name} = {val} in ffi.rs, {hc[name]} in the header"
    for must in ("FDTD_B200_ABI_VERSION", "FDTD_B200_EXACT_ARITHMETIC", "FDTD_B200_IN_PLACE_STEP", "FDTD_B200_OK", "FDTD_B200_EINVAL",
                 "FDTD_B200_ENOMEM", "FDTD_B200_ECUDA", "FDTD_B200_ESTATE", "FDTD_B200_FIELD_E", "FDTD_B200_FIELD_H",
                 "FDTD_B200_COLORMAP_CPU", "FDTD_B200_COLORMAP_LINEAR", "FDTD_B200_COLORMAP_APP_Z", "FDTD_B200_P2P_BLOB_BYTES"):
        assert must in rc, must


TRAIT_IMPLS = [
    # (file, regex) — every trait the runner's bounds name (cem-app/src/solver/runner.rs:247-259, 443-450)
    ("lib.rs", r"impl SolverBackend<FdtdSolverConfig, Point3<usize>> for FdtdB200Backend"),
    ("lib.rs", r"fn create_instance<D>"),
    ("lib.rs", r"fn memory_required\(&self, config: &FdtdSolverConfig\) -> Option<usize>"),
    ("lib.rs", r"impl SolverInstance for FdtdB200SolverInstance"),
    ("lib.rs", r"fn create_state\(&self\) -> FdtdB200SolverState"),
    ("lib.rs", r"fn begin_update<'a>"),
    ("lib.rs", r"impl UpdatePassForcing<Point3<usize>> for FdtdB200UpdatePass<'_>"),
    ("lib.rs", r"impl UpdatePass for FdtdB200UpdatePass<'_>"),
    ("lib.rs", r"impl Time for FdtdB200SolverState"),
    ("lib.rs", r"unsafe impl Send for FdtdB200SolverState"),
    ("lib.rs", r"impl Drop for FdtdB200SolverState"),
    ("field.rs", r"impl Field<Point3<usize>> for FdtdB200SolverInstance"),
    ("field.rs", r"impl FieldView<Point3<usize>> for B200FieldView"),
    ("field.rs", r"fn at\(&self, point: &Point3<usize>\) -> Option<Vector3<f64>>"),
    ("field.rs", r"impl Iterator for B200FieldIter<'_>"),
    ("project.rs", r"impl<Target> CreateProjection<Target> for FdtdB200SolverInstance"),
    ("project.rs", r"impl BeginProjectionPass for FdtdB200SolverInstance"),
    ("project.rs", r"impl<'a, Target> ProjectionPassAdd<'a, B200ImageProjection<Target>> for B200ProjectionPass<'a>"),
    ("project.rs", r"impl ProjectionPass for B200ProjectionPass<'_>"),
    ("project.rs", r"fn finish\(self\) -> Result<\(\), B200ProjectionPassError>"),
    ("pml.rs", r"pub struct PmlProfileBuilder"),
    ("error.rs", r"pub enum B200Error"),
    ("error.rs", r"pub fn check\(rc: c_int\)"),
    ("error.rs", r"pub fn last_error\(h: ffi::fdtd_b200_handle\)"),
]


@pytest.mark.parametrize("fname,pattern", TRAIT_IMPLS)
def test_trait_layer_is_written_out(fname, pattern):
    txt = open(os.path.join(CRATE, "src", fname)).read()
    assert re.search(pattern, txt), f"{fname}: missing `{pattern}`"


def test_crate_files_and_ffi_calls_are_consistent():
    for f in ("Cargo.toml", "build.rs", "src/ffi.rs", "src/lib.rs", "src/error.rs", "src/lattice.rs", "src/pml.rs", "src/field.rs",
              "src/project.rs"):
        assert os.path.exists(os.path.join(CRATE, f)), f
    assert os.path.exists(os.path.join(ROOT, "rust", "cem-app-patch", "runner_b200.rs"))
    rf, _, rc = parse_ffi_rs()
    # every ffi:: item used by the safe layer exists in ffi.rs; no todo!/unimplemented! anywhere
    for f in ("lib.rs", "error.rs", "field.rs", "project.rs", "pml.rs", "lattice.rs"):
        txt = open(os.path.join(CRATE, "src", f)).read()
        assert "todo!" not in txt and "unimplemented!" not in txt, f
        for item in set(re.findall(r"ffi::(fdtd_b200_\w+|FDTD_B200_\w+)", txt)):
            assert item in rf or item in rc or item in ("fdtd_b200_handle", "fdtd_b200_config", "fdtd_b200_material", "fdtd_b200_source",
                                                         "fdtd_b200_physical_constants"), f"{f} uses ffi::{item}, which ffi.rs does not declare"
    patch = open(os.path.join(ROOT, "rust", "cem-app-patch", "runner_b200.rs")).read()
    assert "Parallelization::B200" in patch and "impl CreateProjection<TextureSenderTarget> for FdtdB200SolverInstance" in patch
