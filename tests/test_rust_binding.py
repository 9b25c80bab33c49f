"""The Rust FFI crate (rust/raysnail-b200-sys) cannot be compiled here (no cargo/rustc), so
its #[repr(C)] structs and extern "C" declarations are compared TEXTUALLY with the C header
(include/raysnail_b200.h): same structs, same fields in the same order with matching types,
same functions with the same number of parameters, same constants."""
import os
import re

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HEADER = open(os.path.join(ROOT, "include", "raysnail_b200.h")).read()
RUST = open(os.path.join(ROOT, "rust", "raysnail-b200-sys", "src", "lib.rs")).read()

C2RUST = {"int32_t": "i32", "uint32_t": "u32", "int64_t": "i64", "uint64_t": "u64", "double": "f64", "float": "f32", "uint8_t": "u8", "int": "c_int"}


def strip_comments(text):
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return re.sub(r"//[^\n]*", "", text)


def c_structs():
    out = {}
    for m in re.finditer(r"typedef struct (\w+) \{(.*?)\} \1;", strip_comments(HEADER), flags=re.S):
        fields = []
        for decl in m.group(2).split(";"):
            decl = " ".join(decl.split())
            if not decl:
                continue
            mm = re.match(r"(const )?(\w+)(\*?) ?(.*)$", decl)
            const, ctype, star, names = mm.group(1), mm.group(2), mm.group(3), mm.group(4)
            for name in names.split(","):
                name = name.strip()
                arr = re.match(r"(\w+)\[(\d+)\]", name)
                base = C2RUST.get(ctype, ctype)
                if star:
                    rtype = f"*const {base}" if const else f"*mut {base}"
                    fields.append((name, rtype))
                elif arr:
                    fields.append((arr.group(1), f"[{base}; {arr.group(2)}]"))
                else:
                    fields.append((name, base))
        out[m.group(1)] = fields
    return out


def rust_structs():
    out = {}
    for m in re.finditer(r"pub struct (\w+) \{(.*?)\n\}", strip_comments(RUST), flags=re.S):
        fields = []
        for line in m.group(2).split("\n"):
            mm = re.match(r"\s*pub (\w+): (.+?),\s*$", line)
            if mm:
                fields.append((mm.group(1).rstrip("_") if mm.group(1) == "type_" else mm.group(1), mm.group(2)))
        out[m.group(1)] = fields
    return out


def test_structs_match_the_header_field_for_field():
    c, r = c_structs(), rust_structs()
    names = ["rs_node", "rs_xform", "rs_material", "rs_texture", "rs_scene_desc", "rs_camera", "rs_render_params", "rs_stats"]
    for n in names:
        assert n in c and n in r, n
        assert c[n] == r[n], f"{n}:\n C    {c[n]}\n Rust {r[n]}"


def c_functions():
    text = " ".join(strip_comments(HEADER).split())
    out = {}
    for m in re.finditer(r"(?:int|void|uint64_t|const char\*) (rs_\w+)\((.*?)\);", text):
        args = m.group(2).strip()
        out[m.group(1)] = 0 if args in ("", "void") else len(args.split(","))
    return out


def rust_functions():
    text = " ".join(strip_comments(RUST).split())
    block = re.search(r'extern "C" \{(.*?)\n?\} ', text + " ")
    out = {}
    for m in re.finditer(r"pub fn (rs_\w+)\((.*?)\)(?: -> [^;]+)?;", text):
        args = m.group(2).strip()
        out[m.group(1)] = 0 if not args else len(args.split(","))
    return out


def test_every_entry_point_is_declared_with_the_same_arity():
    c, r = c_functions(), rust_functions()
    assert len(c) >= 20
    assert set(c) == set(r), (sorted(set(c) - set(r)), sorted(set(r) - set(c)))
    for name, n in c.items():
        assert r[name] == n, (name, n, r[name])


def test_constants_match():
    ctext = strip_comments(HEADER)
    cvals = {m.group(1): int(m.group(2)) for m in re.finditer(r"\b(RS_[A-Z0-9_]+) = (\d+)", ctext)}
    cvals.update({m.group(1): int(m.group(2)) for m in re.finditer(r"#define (RS_[A-Z0-9_]+) (\d+)u?\b", ctext)})
    rvals = {m.group(1): int(m.group(2)) for m in re.finditer(r"pub const (RS_[A-Z0-9_]+): \w+ = (\d+);", RUST)}
    assert len(rvals) > 30
    for name, v in rvals.items():
        assert name in cvals, name
        assert cvals[name] == v, (name, cvals[name], v)
    for name in cvals:
        assert name in rvals, f"{name} missing from the Rust crate"
