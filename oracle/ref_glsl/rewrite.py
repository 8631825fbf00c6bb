#!/usr/bin/env python3
"""rewrite.py <reference shader dir> <out dir>

Textual GLSL -> C++ rewrites of the REFERENCE's own shader sources, so that they compile under g++
against oracle/ref_glsl/glsl_compat.h.  Output goes to oracle/_ref/ (git-ignored); reference sources
are never copied into the repo.  Only what C++ syntax cannot express is touched:

  R1  floating-point literals get an `f` suffix (a GLSL literal is fp32, a C++ one is double)
  R2  `inout T name` parameters become `T& name`
  R3  `layout(...)` interface declarations become plain C++ declarations of the same names:
        buffer_reference blocks  -> GLSL_BUFFER_REFERENCE(Name, type, field)  (typed view of an address)
        storage blocks           -> struct Name { const T* field; } instance;
        sampler / image / acceleration-structure uniforms, push constants -> globals of the same name
        specialization constants -> plain globals (set by the harness)
        rayPayloadEXT            -> thread_local global; rayPayloadInEXT -> the enclosing scope's `payload`
        hitAttributeEXT          -> thread_local global
  R4  `#version` / `#extension` lines are dropped

Statements, expressions, identifiers, control flow and evaluation order are left exactly as written.
Every rule reports how many times it fired; a rule that never fires on a file that should need it is an
error, so a changed reference cannot silently slip through.
"""
import os
import re
import sys

FILES = ["rand.glsl", "common.glsl", "triangle.glsl", "material.glsl", "brdf.glsl", "ray_gen.rgen", "closest_hit.rchit", "miss.rmiss"]

LITERAL = re.compile(r"(?<![\w.])(\d+\.\d*(?:[eE][-+]?\d+)?|\d+[eE][-+]?\d+)(?![\w.])")
INOUT = re.compile(r"\binout\s+(\w+)\s+(\w+)")
BUFFER_REF = re.compile(r"layout\(buffer_reference,\s*scalar\)\s*readonly\s+buffer\s+(\w+)\s*\{\s*(\w+)\s+(\w+)\[\];\s*\};")
STORAGE = re.compile(r"layout\(binding\s*=\s*\d+,\s*set\s*=\s*\d+\)\s*readonly\s+buffer\s+(\w+)\s*\{\s*(\w+)\s+(\w+)\[\];\s*\}\s*(\w+);")
SAMPLERS = re.compile(r"layout\(binding\s*=\s*\d+,\s*set\s*=\s*\d+\)\s*uniform\s+sampler2D\s+(\w+)\[\];")
SPEC_CONST = re.compile(r"layout\(constant_id\s*=\s*\d+\)\s*const\s+(\w+)\s+(\w+)\s*=")
OPAQUE = re.compile(r"layout\(binding\s*=\s*\d+,\s*set\s*=\s*\d+\)\s*(?:writeonly\s+)?uniform\s+(accelerationStructureEXT|image2D)\s+(\w+);")
PUSH = re.compile(r"layout\(push_constant\)\s*uniform\s+(\w+)\s*\{")
PAYLOAD = re.compile(r"layout\(location\s*=\s*\d+\)\s*rayPayloadEXT\s+(\w+)\s+(\w+);")
PAYLOAD_IN = re.compile(r"layout\(location\s*=\s*\d+\)\s*rayPayloadInEXT\s+(\w+)\s+(\w+);")
HIT_ATTR = re.compile(r"hitAttributeEXT\s+(\w+)\s+(\w+);")
DIRECTIVE = re.compile(r"^#(version|extension)\b.*$", re.M)

# rule -> files in which it must fire at least once
MUST_FIRE = {
    "literal": {"rand.glsl", "material.glsl", "brdf.glsl", "ray_gen.rgen"},
    "inout": {"rand.glsl"},
    "buffer_reference": {"triangle.glsl"},
    "storage": {"triangle.glsl", "material.glsl"},
    "samplers": {"material.glsl"},
    "spec_const": {"ray_gen.rgen"},
    "opaque": {"ray_gen.rgen"},
    "push": {"ray_gen.rgen"},
    "payload": {"ray_gen.rgen"},
    "payload_in": {"closest_hit.rchit", "miss.rmiss"},
    "hit_attr": {"closest_hit.rchit"},
}


def rewrite(name, src):
    fired = {}

    def sub(rule, rx, repl, text):
        out, n = rx.subn(repl, text)
        fired[rule] = n
        return out

    src = sub("directive", DIRECTIVE, "", src)
    src = sub("literal", LITERAL, r"\1f", src)
    src = sub("inout", INOUT, r"\1& \2", src)
    src = sub("buffer_reference", BUFFER_REF, r"GLSL_BUFFER_REFERENCE(\1, \2, \3)", src)
    src = sub("storage", STORAGE, r"struct \1 { const \2* \3; } \4;", src)
    src = sub("samplers", SAMPLERS, r"const sampler2D* \1;", src)
    src = sub("spec_const", SPEC_CONST, r"\1 \2 =", src)
    src = sub("opaque", OPAQUE, r"\1 \2;", src)
    src = sub("push", PUSH, r"struct \1 {", src)
    src = sub("payload", PAYLOAD, r"thread_local \1 \2;", src)
    src = sub("payload_in", PAYLOAD_IN, r"/* \2: the ray payload of the enclosing scope */", src)
    src = sub("hit_attr", HIT_ATTR, r"thread_local \1 \2;", src)
    for rule, files in MUST_FIRE.items():
        if name in files and not fired.get(rule):
            raise SystemExit(f"rewrite.py: rule {rule} did not fire on {name}: the reference changed, review the rewrite")
    if "layout(" in src or "inout" in src:
        raise SystemExit(f"rewrite.py: unhandled layout()/inout left in {name}")
    return src, fired


def main():
    src_dir, out_dir = sys.argv[1], sys.argv[2]
    os.makedirs(out_dir, exist_ok=True)
    for name in FILES:
        with open(os.path.join(src_dir, name)) as f:
            text = f.read()
        out, fired = rewrite(name, text)
        with open(os.path.join(out_dir, name), "w") as f:
            f.write(out)
        print(f"  {name}: " + ", ".join(f"{k} x{v}" for k, v in fired.items() if v))


if __name__ == "__main__":
    main()
