"""The Rust host side ships as SOURCE (rust/, not compiled on this box: no cargo).  What can be checked without a toolchain is
checked here: the `#[repr(C)]` structs of rust/pathtrace-b200-sys/src/lib.rs against include/pathtrace_b200.h (field names,
order, sizes, offsets under the C layout rules), its `extern "C"` block against the header's exported symbols, its constants
against the header's enums, and that rust/reference-describe.patch applies cleanly to the reference tree where that exists."""
import os
import re
import shutil
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HEADER = open(os.path.join(ROOT, "include", "pathtrace_b200.h")).read()
HEADER_NC = re.sub(r"/\*.*?\*/", "", HEADER, flags=re.S)
RUST = open(os.path.join(ROOT, "rust", "pathtrace-b200-sys", "src", "lib.rs")).read()
RUST_NC = re.sub(r"//.*", "", RUST)

C_SIZES = {"uint32_t": 4, "int32_t": 4, "uint64_t": 8, "float": 4, "uint8_t": 1}
R_SIZES = {"u32": 4, "i32": 4, "u64": 8, "f32": 4, "u8": 1}


def c_struct(name):
    body = re.search(r"typedef struct \{([^{}]*)\}\s*" + name + r"\s*;", HEADER_NC).group(1)
    fields = []
    for decl in body.split(";"):
        decl = decl.strip()
        if not decl:
            continue
        ty, rest = decl.split(None, 1)
        for item in rest.split(","):
            m = re.match(r"\s*(\w+)(?:\[(\d+)\])?\s*$", item)
            fields.append((m.group(1), C_SIZES[ty], int(m.group(2) or 1)))
    return fields


def rust_struct(name):
    body = re.search(r"#\[repr\(C\)\][^{}]*?pub struct " + name + r"\s*\{([^{}]*)\}", RUST_NC, flags=re.S).group(1)
    fields = []
    for m in re.finditer(r"pub (\w+):\s*(\[(\w+);\s*(\d+)\]|\w+)", body):
        ty = m.group(3) or m.group(2)
        fields.append((m.group(1), R_SIZES[ty], int(m.group(4) or 1)))
    return fields


def layout(fields):
    off, align, out = 0, 1, []
    for name, sz, n in fields:
        off = (off + sz - 1) // sz * sz
        out.append((name, off, sz * n))
        off += sz * n
        align = max(align, sz)
    return out, (off + align - 1) // align * align


@pytest.mark.parametrize("name", ["pt_shape", "pt_material", "pt_camera", "pt_render_params", "pt_stats"])
def test_repr_c_structs_match_header(name):
    c, r = c_struct(name), rust_struct(name)
    assert [f[0] for f in c] == [f[0] for f in r], name
    assert layout(c) == layout(r), name
    assert "#[repr(C)]" in RUST


def test_struct_sizes_are_the_documented_ones():
    assert layout(rust_struct("pt_shape"))[1] == 40 and layout(rust_struct("pt_material"))[1] == 36
    assert layout(rust_struct("pt_camera"))[1] == 56 and layout(rust_struct("pt_render_params"))[1] == 48


def test_extern_block_declares_every_export():
    declared = set(re.findall(r"\b(pt_[a-z0-9_]+)\s*\(", HEADER_NC))
    rust_fns = set(re.findall(r"pub fn (pt_[a-z0-9_]+)\s*\(", RUST_NC))
    assert declared == rust_fns, declared ^ rust_fns


def test_constants_match_header_enums():
    enums = dict((k, int(v)) for k, v in re.findall(r"\b(PT_[A-Z0-9_]+)\s*=\s*(\d+)", HEADER_NC))
    enums.update((k, int(v)) for k, v in re.findall(r"#define\s+(PT_[A-Z0-9_]+)\s+(\d+)u?\b", HEADER_NC))
    consts = dict((k, int(v)) for k, v in re.findall(r"pub const (PT_[A-Z0-9_]+):\s*\w+\s*=\s*(\d+);", RUST_NC))
    assert consts, "no constants parsed"
    for k, v in consts.items():
        assert enums.get(k) == v, (k, v, enums.get(k))
    for k in ("PT_TRAVERSAL_BRUTE_EXACT", "PT_FLAG_COUNTERS", "PT_MAT_OREN_NAYAR", "PT_STRATEGY_BRDF_ONLY"):
        assert k in consts


def test_wrapper_fills_every_field_of_render_params():
    wrapper = open(os.path.join(ROOT, "rust", "pathtrace-b200", "src", "lib.rs")).read()
    lit = re.search(r"sys::pt_render_params \{(.*?)\}", wrapper, flags=re.S).group(1)
    for name, _, _ in c_struct("pt_render_params"):
        assert re.search(r"\b" + name + r"\s*:", lit), name


def test_reference_patch_applies(tmp_path):
    ref = "/root/reference"
    if not os.path.isdir(ref) or shutil.which("patch") is None:
        pytest.skip("reference tree (or patch) not available on this machine")
    dst = os.path.join(tmp_path, "ref")
    shutil.copytree(ref, dst)
    subprocess.check_call(["chmod", "-R", "u+w", dst])
    r = subprocess.run(["patch", "-p1", "--dry-run", "-i", os.path.join(ROOT, "rust", "reference-describe.patch")], cwd=dst, capture_output=True, text=True)
    assert r.returncode == 0, r.stdout + r.stderr
    assert "src/main.rs" in r.stdout and "src/objects/shape.rs" in r.stdout
