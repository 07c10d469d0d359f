"""The Rust side of the drop-in (rust/) cannot be compiled here (no cargo / rustc in this image), so it is checked
against the C header instead: the generated extern block declares every RT_API symbol with the same arity and the same
pointer / integer widths, the committed file is what the generator produces from the current header, and the
hand-written wrapper and node only call functions that exist."""
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "tools"))
import gen_rust_bindings as gen  # noqa: E402

WIDTH_C = {"int": 4, "uint32_t": 4, "uint64_t": 8, "uint8_t": 1, "float": 4, "double": 8, "char": 1}
WIDTH_RS = {"c_int": 4, "i32": 4, "u32": 4, "u64": 8, "u8": 1, "f32": 4, "f64": 8, "c_char": 1}


def c_shape(ctype):
    """(pointer depth, width in bytes of the pointee / value or the struct name)."""
    toks = [t for t in ctype.replace("*", " * ").split() if t not in ("const", "struct")]
    base, depth = toks[0], toks.count("*")
    return depth, WIDTH_C.get(base, base)


def rs_shape(rtype):
    depth = rtype.count("*")
    base = re.sub(r"\*(const|mut)\s+", "", rtype).strip().split("::")[-1]
    return depth, WIDTH_RS.get(base, "void" if base == "c_void" else base)


def rust_functions():
    text = open(os.path.join(ROOT, "rust", "src", "rt_b200.rs")).read()
    block = text[text.index('extern "C" {'):]
    out = {}
    for m in re.finditer(r"pub fn (rt_\w+)\(([^)]*)\)(?:\s*->\s*([^;]+))?;", block):
        params = [p.split(":", 1)[1].strip() for p in m.group(2).split(",") if p.strip()]
        out[m.group(1)] = (params, (m.group(3) or "").strip())
    return out


def test_generated_file_is_current():
    assert subprocess.run([sys.executable, os.path.join(ROOT, "tools", "gen_rust_bindings.py"), "--check"]).returncode == 0


def test_extern_block_matches_header():
    header = gen.strip_comments(open(gen.HEADER).read())
    c_funcs = {name: (ret, params) for name, ret, params in gen.parse_functions(header)}
    rs_funcs = rust_functions()
    assert len(c_funcs) >= 60
    assert sorted(c_funcs) == sorted(rs_funcs)
    for name, (ret, params) in c_funcs.items():
        rs_params, rs_ret = rs_funcs[name]
        assert len(params) == len(rs_params), name
        for (ctype, pname), rtype in zip(params, rs_params):
            cd, cw = c_shape(ctype)
            rd, rw = rs_shape(rtype)
            assert cd == rd and str(cw) == str(rw), (name, pname, ctype, rtype)
            if cd >= 1:   # constness of the first pointer level
                assert ("const" in ctype.split("*")[0]) == rtype.startswith("*const"), (name, pname, ctype, rtype)
        if ret == "void":
            assert rs_ret == ""
        else:
            assert c_shape(ret) == rs_shape(rs_ret) or (c_shape(ret)[0] == rs_shape(rs_ret)[0] and str(c_shape(ret)[1]) == str(rs_shape(rs_ret)[1])), name


def test_wrapper_and_node_call_declared_functions_only():
    declared = set(rust_functions())
    used = set()
    for f in ("renderer.rs", "node.rs"):
        used |= set(re.findall(r"ffi::(rt_\w+)", open(os.path.join(ROOT, "rust", "src", f)).read()))
    assert used and used <= declared, used - declared
    # the constants the wrapper names exist too
    text = open(os.path.join(ROOT, "rust", "src", "rt_b200.rs")).read()
    for const in re.findall(r"ffi::(RT_\w+)", open(os.path.join(ROOT, "rust", "src", "renderer.rs")).read()):
        assert f"pub const {const}:" in text, const


def test_struct_layouts_match_the_header():
    """CameraSettings / RtVec3 as generated: field order and widths of the C structs (40 + 12-byte vectors)."""
    header = gen.strip_comments(open(gen.HEADER).read())
    text = open(os.path.join(ROOT, "rust", "src", "rt_b200.rs")).read()
    for sname in gen.LOCAL_STRUCTS:
        body = re.search(r"pub struct " + sname + r" \{(.*?)\n\}", text, flags=re.S).group(1)
        rs_fields = re.findall(r"pub (\w+): (\w+),", body)
        c_fields = gen.parse_struct(header, sname)
        assert [n for n, _ in rs_fields] == [n for _, n in c_fields]
        for (rn, rtype), (ctype, cn) in zip(rs_fields, c_fields):
            assert str(WIDTH_RS.get(rtype, rtype)) == str(WIDTH_C.get(ctype, ctype)), (sname, rn)
