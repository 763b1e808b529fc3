"""The Rust binding (rust/datafusion-gpu/src/sys.rs) against include/dfgpu.h, mechanically: every #[repr(C)] struct has the
header's fields in the header's order with the matching types, every function the header declares is declared in the `extern "C"`
block with the same parameter list, and the enum constants agree.  There is no Rust toolchain in this image, so this is what
keeps the two texts from drifting (round 1's sketch lacked dfgpu_options.coalesce_rows: dfgpu_options_default would have written
8 bytes past the Rust struct)."""
import os
import re

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HDR = open(os.path.join(ROOT, "include", "dfgpu.h")).read()
SYS = open(os.path.join(ROOT, "rust", "datafusion-gpu", "src", "sys.rs")).read()
HDR_NC = re.sub(r"/\*.*?\*/", "", HDR, flags=re.S)
SYS_NC = re.sub(r"//[^\n]*", "", SYS)

BASE = {"int32_t": "i32", "int64_t": "i64", "uint64_t": "u64", "uint8_t": "u8", "double": "f64", "char": "c_char", "void": "c_void",
        "dfgpu_status": "c_int", "struct ArrowArray": "FFI_ArrowArray", "struct ArrowSchema": "FFI_ArrowSchema"}


def rust_type(c):
    """C declarator type (without the name) -> the Rust FFI type it must be bound as."""
    c = " ".join(c.split())
    m = re.match(r"^(const )?(.*?)\s*((?:\*\s*(?:const\s*)?)*)$", c)
    const0, core, ptrs = bool(m.group(1)), m.group(2).strip(), m.group(3)
    out = BASE.get(core, core)
    levels = re.findall(r"\*\s*(const)?", ptrs)
    for i in range(len(levels)):
        is_const = const0 if i == 0 else levels[i - 1] == "const"
        out = ("*const " if is_const else "*mut ") + out
    return out


def header_structs():
    out = {}
    for m in re.finditer(r"typedef struct (\w+) \{(.*?)\} \1;", HDR_NC, flags=re.S):
        fields = []
        for decl in m.group(2).split(";"):
            decl = " ".join(decl.split())
            if not decl:
                continue
            mm = re.match(r"^(.*?)([\w, ]+)$", decl)
            typ, names = mm.group(1), [n.strip() for n in mm.group(2).split(",")]
            # "int32_t rank, world" declares several fields; "const void *nccl_unique_id" keeps its pointer with the type
            first = re.match(r"^(.*?)(\w+)$", (typ + names[0]).strip())
            base_t = first.group(1)
            fields.append((first.group(2), rust_type(base_t)))
            for n in names[1:]:
                fields.append((n, rust_type(re.sub(r"\*\s*$", "", base_t))))
        out[m.group(1)] = fields
    return out


def rust_structs():
    out = {}
    for m in re.finditer(r"#\[repr\(C\)\]\s*(?:#\[derive\([^\)]*\)\]\s*)?pub struct (\w+)\s*\{(.*?)\}", SYS_NC, flags=re.S):
        fields = [(f.group(1), " ".join(f.group(2).split())) for f in re.finditer(r"pub (\w+):\s*([^,\n]+),?", m.group(2))]
        out[m.group(1)] = fields
    return out


def header_functions():
    out = {}
    for m in re.finditer(r"^([a-zA-Z_][\w \*]*?)\b(dfgpu_\w+)\s*\(([^;]*?)\)\s*;", HDR_NC, flags=re.M | re.S):
        ret, name, params = m.group(1).strip(), m.group(2), " ".join(m.group(3).split())
        ps = []
        if params != "void":
            for p in params.split(","):
                mm = re.match(r"^(.*?)(\w+)$", p.strip())
                ps.append((mm.group(2), rust_type(mm.group(1))))
        out[name] = (ps, None if ret == "void" else rust_type(ret))
    return out


def rust_functions():
    block = re.search(r'extern "C" \{(.*?)\n\}', SYS_NC, flags=re.S).group(1)
    out = {}
    for m in re.finditer(r"pub fn (\w+)\s*\((.*?)\)\s*(?:->\s*([^;]+))?;", block, flags=re.S):
        ps = [(p.group(1), " ".join(p.group(2).split())) for p in re.finditer(r"(\w+):\s*([^,\)]+?)\s*(?:,|$)", " ".join(m.group(2).split()))]
        out[m.group(1)] = (ps, m.group(3).strip() if m.group(3) else None)
    return out


def test_structs_match_field_by_field():
    h, r = header_structs(), rust_structs()
    assert set(h) == {"dfgpu_options", "dfgpu_metrics", "dfgpu_expr_node", "dfgpu_agg_desc"}
    for name, fields in h.items():
        assert name in r, f"sys.rs lacks #[repr(C)] struct {name}"
        assert r[name] == fields, f"{name}: header {fields} != sys.rs {r[name]}"


def test_every_function_is_bound_with_the_same_signature():
    h, r = header_functions(), rust_functions()
    assert len(h) >= 57
    assert set(h) == set(r), set(h) ^ set(r)
    for name, sig in h.items():
        assert r[name] == sig, f"{name}: header {sig} != sys.rs {r[name]}"


def test_opaque_handles_and_constants():
    for hnd in re.findall(r"typedef struct (\w+) \1;", HDR_NC):
        assert re.search(r"#\[repr\(C\)\]\s*pub struct %s \{ _p: \[u8; 0\] \}" % hnd, SYS), hnd
    consts = {}
    for m in re.finditer(r"enum\s*\{(.*?)\};", HDR_NC, flags=re.S):
        for name, val in re.findall(r"(DFGPU_\w+)\s*=\s*(-?\d+)", m.group(1)):
            consts[name] = int(val)
    assert len(consts) >= 35
    for name, val in consts.items():
        m = re.search(r"pub const %s: \w+ = (-?\d+);" % name, SYS)
        assert m and int(m.group(1)) == val, name
    assert re.search(r"pub const DFGPU_NOALLOC_MAX_ROWS: i64 = 1 << 20;", SYS) and "#define DFGPU_NOALLOC_MAX_ROWS (1 << 20)" in HDR
    assert re.search(r"pub const DFGPU_NOALLOC_MAX_PARTS: i32 = 256;", SYS) and "#define DFGPU_NOALLOC_MAX_PARTS 256" in HDR


def test_shim_never_post_filters_and_guards_collect_left():
    """The two semantic mistakes of the round-1 sketch stay fixed: a residual JoinFilter goes through probe_candidates / probe_commit
    (never applied to the joined output), and CollectLeft is refused for the join types that emit unmatched build rows per partition."""
    hj = open(os.path.join(ROOT, "rust", "datafusion-gpu", "src", "hash_join.rs")).read()
    assert "dfgpu_join_probe_candidates" in hj and "dfgpu_join_probe_commit" in hj and "dfgpu_join_set_filter" in hj
    assert re.search(r"PartitionMode::CollectLeft\s*&&\s*matches!\(cpu\.join_type\(\), JoinType::Left \| JoinType::Full \| JoinType::LeftAnti", hj)
    rule = open(os.path.join(ROOT, "rust", "datafusion-gpu", "src", "rule.rs")).read()
    assert "hash_contract" in rule and "replace both or neither" in rule
