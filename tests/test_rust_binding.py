"""The Rust side of the boundary (bindings/rust) cannot be compiled here (no rustc in the image);
this keeps it honest the only way available: text against text.  Every product entry point that
include/sfp.h declares must be declared in the extern block with the same number of parameters,
and every POD struct must list the same field names in the same order."""
import os
import re

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _strip_comments(text):
    text = re.sub(r"/\*.*?\*/", " ", text, flags=re.S)
    return re.sub(r"//[^\n]*", " ", text)


def header_functions():
    text = _strip_comments(open(os.path.join(ROOT, "include", "sfp.h")).read())
    out = {}
    for m in re.finditer(r"SFP_API\s+[\w\s\*]+?\b(sfp_\w+)\s*\(([^)]*)\)\s*;", text):
        name, args = m.group(1), m.group(2).strip()
        if name.startswith("sfp_ref_"):
            continue
        out[name] = 0 if args in ("", "void") else len(args.split(","))
    return out


def rust_functions():
    text = _strip_comments(open(os.path.join(ROOT, "bindings", "rust", "src", "lib.rs")).read())
    block = text[text.index('extern "C" {'):]
    block = block[:block.index("\n}\n")]
    out = {}
    for m in re.finditer(r"pub fn (sfp_\w+)\s*\(([^)]*)\)", block):
        args = m.group(2).strip()
        out[m.group(1)] = 0 if not args else len([a for a in args.split(",") if a.strip()])
    return out


def header_structs():
    text = _strip_comments(open(os.path.join(ROOT, "include", "sfp.h")).read())
    out = {}
    for m in re.finditer(r"typedef struct (\w+)\s*\{(.*?)\}\s*\1\s*;", text, flags=re.S):
        fields = []
        for decl in m.group(2).split(";"):
            decl = decl.strip()
            if not decl:
                continue
            names = decl.split(None, 1)[1] if " " in decl else decl
            for n in names.split(","):
                fields.append(re.sub(r"\[.*?\]", "", n).strip())
        out[m.group(1)] = fields
    return out


def rust_structs():
    text = _strip_comments(open(os.path.join(ROOT, "bindings", "rust", "src", "lib.rs")).read())
    out = {}
    for m in re.finditer(r"pub struct (\w+)\s*\{(.*?)\}", text, flags=re.S):
        out[m.group(1)] = re.findall(r"pub (\w+)\s*:", m.group(2))
    return out


def test_every_entry_point_is_bound_with_the_same_arity():
    h, r = header_functions(), rust_functions()
    assert len(h) >= 35
    assert set(h) == set(r), (sorted(set(h) - set(r)), sorted(set(r) - set(h)))
    for name, n in h.items():
        assert r[name] == n, (name, n, r[name])


def test_every_pod_has_the_same_fields():
    h, r = header_structs(), rust_structs()
    assert {"SfpParams", "SfpBodyDesc", "SfpJointDesc", "SfpBodyState", "SfpPolygon", "SfpPair",
            "SfpContact", "SfpContactEvent", "SfpRay", "SfpRayHit", "SfpStats"} <= set(h)
    for name, fields in h.items():
        assert name in r, name
        assert r[name] == fields, (name, fields, r[name])


def test_build_rs_keeps_the_flags_that_parity_depends_on():
    text = open(os.path.join(ROOT, "bindings", "rust", "build.rs")).read()
    mk = open(os.path.join(ROOT, "starframe_b200", "csrc", "Makefile")).read()
    for flag in ("--fmad=false", "arch=compute_100a,code=sm_100a", "-ffp-contract=off"):
        assert flag in text and flag in mk, flag
