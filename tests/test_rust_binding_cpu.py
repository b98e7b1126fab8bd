"""CPU: the Rust `-sys` crate under rust/ (source only -- no Rust toolchain here, see rust/README.md)
is kept consistent with the C ABI mechanically: every `extern "C"` function, `#[repr(C)]` struct
and constant of rust/rt_b200-sys/src/lib.rs is checked against include/rt_b200.h,
include/rt_b200_host.h and the ctypes mirror the parity tests call through (rust_tracer_b200/abi.py)."""
import ctypes as C
import os
import re

from conftest import ROOT
from rust_tracer_b200 import _lib, abi

RS = open(os.path.join(ROOT, "rust", "rt_b200-sys", "src", "lib.rs")).read()
RS_NOCOMMENT = re.sub(r"//[^\n]*", "", RS)


def _c_prototypes():
    protos = {}
    for header in ("rt_b200.h", "rt_b200_host.h"):
        text = open(os.path.join(ROOT, "include", header)).read()
        text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
        for name, args in re.findall(r"\b(rt_[a-z0-9_]+)\s*\(([^)]*)\)\s*;", text):
            args = args.strip()
            protos[name] = [] if args in ("", "void") else [a.strip() for a in args.split(",")]
    return protos


def _rust_functions():
    block = re.search(r'extern "C" \{(.*?)\n\}', RS_NOCOMMENT, flags=re.S).group(1)
    fns = {}
    for name, args, ret in re.findall(r"pub fn (rt_[a-z0-9_]+)\s*\((.*?)\)\s*(->\s*[^;]+)?;", block, flags=re.S):
        args = [a.strip() for a in args.split(",") if a.strip()]
        fns[name] = (args, ret.replace("->", "").strip())
    return fns


def test_every_c_function_has_a_rust_declaration_with_the_same_arity():
    c, rs = _c_prototypes(), _rust_functions()
    assert set(c) == set(_lib.EXPORTS)
    assert set(rs) == set(c), set(rs) ^ set(c)
    for name, cargs in c.items():
        assert len(rs[name][0]) == len(cargs), (name, rs[name][0], cargs)


# C parameter type -> Rust parameter type
_TYPES = [
    (r"^const char \*", "*const c_char"), (r"^const float \*", "*const f32"), (r"^float \*", "*mut f32"),
    (r"^const int32_t \*", "*const i32"), (r"^int32_t \*", "*mut i32"), (r"^uint64_t \*", "*mut u64"),
    (r"^uint8_t \*", "*mut u8"), (r"^const uint8_t ", "*const u8"), (r"^uint8_t \w+\[", "*mut u8"),
    (r"^double \*", "*mut f64"), (r"^void \*\*", "*mut *mut c_void"), (r"^void \*", "*mut c_void"),
    (r"^const (rt_\w+) \*", r"*const \1"), (r"^(rt_\w+) \*\*", r"*mut *mut \1"), (r"^(rt_\w+) \*", r"*mut \1"),
    (r"^int32_t ", "i32"), (r"^uint32_t ", "u32"), (r"^uint64_t ", "u64"), (r"^size_t ", "usize"), (r"^float ", "f32"),
]
_RETURNS = {"int": "c_int", "int32_t": "i32", "void": "", "const char *": "*const c_char"}


def _rust_type_of(c_arg):
    for pat, rust in _TYPES:
        m = re.match(pat, c_arg)
        if m:
            return m.expand(rust) if "\\1" in rust else rust
    raise AssertionError("no Rust type known for C parameter %r" % c_arg)


def test_parameter_and_return_types_match():
    c, rs = _c_prototypes(), _rust_functions()
    headers = "".join(re.sub(r"/\*.*?\*/", "", open(os.path.join(ROOT, "include", h)).read(), flags=re.S)
                      for h in ("rt_b200.h", "rt_b200_host.h"))
    for name, cargs in c.items():
        for c_arg, r_arg in zip(cargs, rs[name][0]):
            assert r_arg.split(":", 1)[1].strip() == _rust_type_of(c_arg), (name, c_arg, r_arg)
        ret = re.search(r"([\w ]+?\*?)\s*\b%s\s*\(" % name, headers).group(1).strip()
        assert rs[name][1] == _RETURNS[ret], (name, ret, rs[name][1])


_SIZES = {"i32": (4, 4), "u32": (4, 4), "f32": (4, 4), "u64": (8, 8), "f64": (8, 8)}


def _rust_structs():
    out = {}
    for name, body in re.findall(r"#\[repr\(C\)\][^{]*?pub struct (rt_\w+)\s*\{(.*?)\}", RS_NOCOMMENT, flags=re.S):
        out[name] = [(f, t.strip()) for f, t in re.findall(r"pub (\w+):\s*([^,\n]+),", body)]
    return out


def _layout(fields, structs):
    """(size, align) of a repr(C) struct from its Rust field types."""
    off, align = 0, 1
    for _, t in fields:
        m = re.match(r"\[(\w+); (\d+)\]", t)
        if m:
            size, a = _type_layout(m.group(1), structs)
            size *= int(m.group(2))
        else:
            size, a = _type_layout(t, structs)
        off = (off + a - 1) // a * a + size
        align = max(align, a)
    return (off + align - 1) // align * align, align


def _type_layout(t, structs):
    if t.startswith("*"):
        return 8, 8
    if t in _SIZES:
        return _SIZES[t]
    return _layout(structs[t], structs)


def test_structs_have_the_fields_and_sizes_of_the_ctypes_mirror():
    structs = _rust_structs()
    for name in ("rt_shape", "rt_colorfn", "rt_material", "rt_light", "rt_scene_desc", "rt_camera", "rt_render_params",
                 "rt_stats", "rt_forest_stats"):
        mirror = getattr(abi, name)
        assert [f for f, _ in structs[name]] == [f for f, _ in mirror._fields_], name
        assert _layout(structs[name], structs)[0] == C.sizeof(mirror), name


def test_constants_match():
    consts = {k: int(v) for k, v in re.findall(r"pub const (RT_\w+): \w+ = (-?\d+);", RS_NOCOMMENT)}
    assert consts.pop("RT_B200_ABI_VERSION") == 4
    for k, v in consts.items():
        assert getattr(abi, k) == v, k
    for k in dir(abi):
        if k.startswith("RT_"):
            assert k in consts, k


def test_safe_wrapper_and_patch_only_call_declared_symbols():
    declared = set(_rust_functions())
    for rel in (("rt_b200", "src", "lib.rs"),):
        text = open(os.path.join(ROOT, "rust", *rel)).read()
        used = set(re.findall(r"sys::(rt_[a-z0-9_]+)\s*\(", text))
        assert used and used <= declared, used - declared
