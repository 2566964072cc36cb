"""Pins rust/rtb200-sys/src/lib.rs to include/rtb200.h mechanically (the Rust binding cannot be compiled
in this image — no cargo/rustc — so nothing else would notice it drifting).  Every `pub fn` of the extern block must be
declared in the header with the same arity, argument widths and return width; every header function must be bound; the
#[repr(C)] structs must have the header's fields in the header's order and the sizes the C compiler gives them."""
import ctypes as C
import re
import subprocess
from pathlib import Path

import pytest

ROOT = Path(__file__).resolve().parent.parent
HEADER = ROOT / "include" / "rtb200.h"
FORMAT = ROOT / "include" / "rtb_scene_format.h"
RUST = ROOT / "rust" / "rtb200-sys" / "src" / "lib.rs"

C_SCALARS = {"int": "i32", "int32_t": "i32", "unsigned": "u32", "uint32_t": "u32", "uint64_t": "u64", "size_t": "usize", "double": "f64",
             "float": "f32", "uint8_t": "u8", "char": "i8", "void": "void"}
RUST_SCALARS = {"c_int": "i32", "i32": "i32", "u32": "u32", "u64": "u64", "usize": "usize", "c_double": "f64", "f64": "f64", "f32": "f32",
                "u8": "u8", "c_char": "i8", "c_void": "void"}
SIZES = {"i32": (4, 4), "u32": (4, 4), "u64": (8, 8), "usize": (8, 8), "f64": (8, 8), "f32": (4, 4), "u8": (1, 1), "i8": (1, 1), "ptr": (8, 8)}


def strip_c_comments(text):
    return re.sub(r"/\*.*?\*/", " ", text, flags=re.S)


def c_type_class(t):
    t = t.strip()
    if "(*" in t or "*" in t or t.endswith("]") or t == "rtb_progress_fn":
        return "ptr"
    t = re.sub(r"\b(const|struct|enum)\b", " ", t).split()
    if not t:
        return "void"
    base = t[0]
    if base in C_SCALARS:
        return C_SCALARS[base]
    return "i32" if base.startswith("Rtb") and not base.endswith(("Camera", "Rec", "Table", "Header")) else base  # enums are ints


def parse_c_functions(text):
    text = strip_c_comments(text)
    text = re.sub(r"typedef\s+struct\s+\w*\s*\{.*?\}\s*\w+\s*;", " ", text, flags=re.S)
    text = re.sub(r"typedef\s+enum\s+\w*\s*\{.*?\}\s*\w+\s*;", " ", text, flags=re.S)
    text = re.sub(r"typedef[^;]*\(\s*\*\s*\w+\s*\)\s*\([^;]*\)\s*;", " ", text, flags=re.S)
    text = re.sub(r"^\s*#.*$", " ", text, flags=re.M)
    out = {}
    for m in re.finditer(r"([A-Za-z_][\w\s\*]*?)\b(rtb_\w+)\s*\(([^;{}]*)\)\s*;", text, flags=re.S):
        ret, name, args = m.group(1).strip(), m.group(2), " ".join(m.group(3).split())
        params = [] if args in ("", "void") else [a.strip() for a in args.split(",")]
        classes = []
        for a in params:
            a = re.sub(r"\[[^\]]*\]", "*", a)                       # array parameter = pointer
            names = re.findall(r"[A-Za-z_]\w*", a)
            if "*" not in a and len(names) >= 2 and names[-1] not in C_SCALARS:
                a = a[: a.rfind(names[-1])]                          # drop the parameter name
            classes.append(c_type_class(a))
        out[name] = (c_type_class(ret) if "*" not in ret else "ptr", classes)
    return out


def rust_type_class(t):
    t = t.strip()
    if t.startswith(("*const", "*mut")) or t.startswith("Option<") or t == "rtb_progress_fn":
        return "ptr"
    return RUST_SCALARS.get(t, t)


def parse_rust_functions(text):
    text = re.sub(r"//.*$", "", text, flags=re.M)
    out = {}
    for m in re.finditer(r"pub fn (rtb_\w+)\s*\(([^;]*?)\)\s*(?:->\s*([^;]+?))?\s*;", text, flags=re.S):
        name, args, ret = m.group(1), " ".join(m.group(2).split()), (m.group(3) or "").strip()
        params = [a.split(":", 1)[1].strip() for a in args.split(",") if ":" in a]
        out[name] = ("void" if not ret else rust_type_class(ret), [rust_type_class(p) for p in params])
    return out


def parse_rust_structs(text):
    text = re.sub(r"//.*$", "", text, flags=re.M)
    out = {}
    for m in re.finditer(r"#\[repr\(C\)\]\s*(?:#\[[^\]]*\]\s*)*pub struct (\w+)\s*\{(.*?)\}", text, flags=re.S):
        fields = []
        for f in re.finditer(r"(?:pub\s+)?(\w+)\s*:\s*([^,]+),", m.group(2)):
            fields.append((f.group(1), f.group(2).strip()))
        out[m.group(1)] = fields
    return out


def parse_c_structs(text):
    text = strip_c_comments(text)
    out = {}
    for m in re.finditer(r"typedef\s+struct\s+\w*\s*\{(.*?)\}\s*(\w+)\s*;", text, flags=re.S):
        fields = []
        for decl in m.group(1).split(";"):
            decl = " ".join(decl.split())
            if not decl:
                continue
            ty, rest = decl.split(" ", 1)
            for part in rest.split(","):
                part = part.strip()
                fm = re.match(r"(\w+)((?:\[\d+\])*)$", part)
                assert fm, decl
                dims = [int(x) for x in re.findall(r"\[(\d+)\]", fm.group(2))]
                n = 1
                for d in dims:
                    n *= d
                fields.append((fm.group(1), C_SCALARS[ty], n))
        out[m.group(2)] = fields
    return out


def rust_field(ty):
    m = re.match(r"\[\s*(?:\[\s*(\w+)\s*;\s*(\d+)\s*\]|(\w+))\s*;\s*(\d+)\s*\]$", ty)
    if m:
        inner = m.group(1) or m.group(3)
        n = int(m.group(4)) * (int(m.group(2)) if m.group(2) else 1)
        return RUST_SCALARS[inner], n
    return rust_type_class(ty), 1


def layout_size(fields):
    off, align = 0, 1
    for _, cls, n in fields:
        sz, al = SIZES[cls]
        off = (off + al - 1) // al * al
        off += sz * n
        align = max(align, al)
    return (off + align - 1) // align * align


def test_every_rust_extern_fn_matches_the_header():
    c = parse_c_functions(HEADER.read_text())
    r = parse_rust_functions(RUST.read_text())
    assert len(c) >= 60 and len(r) >= 50
    missing = sorted(set(c) - set(r))
    assert not missing, f"header functions without a Rust binding: {missing}"
    extra = sorted(set(r) - set(c))
    assert not extra, f"Rust bindings with no header declaration: {extra}"
    for name, (ret, args) in r.items():
        cret, cargs = c[name]
        assert len(args) == len(cargs), (name, args, cargs)
        assert ret == cret, (name, "return", ret, cret)
        for i, (a, b) in enumerate(zip(args, cargs)):
            assert a == b, (name, f"argument {i}", a, b)


def test_rust_repr_c_structs_match_the_header_layout(rtb):
    cs = parse_c_structs(HEADER.read_text())
    cs.update(parse_c_structs(FORMAT.read_text()))
    rs = parse_rust_structs(RUST.read_text())
    mirrors = {"RtbCamera": rtb.RtbCamera, "rtb_render_opts": rtb.RenderOpts, "rtb_stats": rtb.Stats}
    for name in ("RtbCamera", "rtb_render_opts", "rtb_stats"):
        assert name in rs and name in cs, name
        rf = [(n,) + rust_field(t) for n, t in rs[name]]
        assert [(n, c, k) for n, c, k in rf] == cs[name], (name, rf, cs[name])
        assert layout_size(cs[name]) == C.sizeof(mirrors[name]), name  # = tests/test_cabi.py::test_struct_layouts_match_ctypes


def test_c_compiler_agrees_with_the_computed_sizes(tmp_path):
    cs = parse_c_structs(HEADER.read_text())
    src = tmp_path / "sz.c"
    src.write_text('#include <stdio.h>\n#include "rtb200.h"\nint main(void){printf("%zu %zu %zu\\n", sizeof(RtbCamera), sizeof(rtb_render_opts), sizeof(rtb_stats));return 0;}\n')
    exe = tmp_path / "sz"
    subprocess.run(["gcc", "-std=c99", f"-I{ROOT / 'include'}", str(src), "-o", str(exe)], check=True)
    got = [int(x) for x in subprocess.run([str(exe)], capture_output=True, text=True, check=True).stdout.split()]
    cs.update(parse_c_structs(FORMAT.read_text()))
    assert got == [layout_size(cs["RtbCamera"]), layout_size(cs["rtb_render_opts"]), layout_size(cs["rtb_stats"])]
