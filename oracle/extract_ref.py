#!/usr/bin/env python3
"""Build-time extraction of reference code fragments into oracle/_ref/gen/ (git-ignored, never committed).

The reference's device header shaders/tracer.cuh cannot be included as a whole (trace()/traceMesh() call OptiX
intrinsics, and two expressions only build under MSVC), but its PURE functions compile as host C++ against
oracle/ref_stub/optix.h.  This script copies those functions -- found by NAME, with brace matching, from the file where it
lies under /root/reference -- into gen/tracer_pure.inc, and three statement ranges of src/GaussianTracer.cpp (the
proxy transform of createGaussianParticlesBVH, the 3x4 packing of createIAS, the normal transform of
sendGeometryAttributesToDevice) into gen/*.inc.  ref_math_shim.cpp includes them.  TEST INFRASTRUCTURE ONLY.

usage: extract_ref.py /root/reference oracle/_ref/gen
"""
import os
import re
import sys

PURE_FUNCS = ["quantizeUnsigned8Bits", "getRay", "getFishEyeRay", "computeResponse", "SHToRadiance", "computeRadiance",
              "renderMirror", "refract"]


def grab_function(src: str, name: str) -> str:
    """text of the definition of `name` (from its storage-class line to the matching closing brace)"""
    pat = r"^" + re.escape(name) + r"\s*$" if name.startswith("struct ") else r"^[^\n;{}]*\b" + re.escape(name) + r"\s*\("
    m = re.search(pat, src, re.M)
    if not m:
        raise SystemExit(f"extract_ref: function {name} not found")
    start = m.start()
    i = src.index("{", m.end())
    depth = 0
    while True:
        c = src[i]
        if c == "{":
            depth += 1
        elif c == "}":
            depth -= 1
            if depth == 0:
                break
        i += 1
    return src[start:i + 1]


def grab_between(src: str, first_pat: str, last_pat: str) -> str:
    """whole lines from the one matching first_pat to the one matching last_pat (inclusive)"""
    lines = src.split("\n")
    a = next(i for i, l in enumerate(lines) if re.search(first_pat, l))
    b = next(i for i in range(a, len(lines)) if re.search(last_pat, lines[i]))
    return "\n".join(lines[a:b + 1])


def main():
    ref, out = sys.argv[1], sys.argv[2]
    os.makedirs(out, exist_ok=True)
    cuh = open(os.path.join(ref, "shaders/tracer.cuh")).read()
    parts = ["// GENERATED from shaders/tracer.cuh by oracle/extract_ref.py -- do not commit\n"]
    consts = [l for l in cuh.split("\n") if re.match(r"\s*(static\s+)?constexpr\s", l)]
    parts.append("\n".join(consts) + "\n")
    parts.append(grab_function(cuh, "struct HitPayload") + ";\n")
    for f in PURE_FUNCS:
        parts.append(grab_function(cuh, f) + "\n")
    open(os.path.join(out, "tracer_pure.inc"), "w").write("\n".join(parts))

    gt = open(os.path.join(ref, "src/GaussianTracer.cpp")).read()
    # createGaussianParticlesBVH: from `float s = sqrtf(` to `glm::mat4 transform = ...;`  (src/GaussianTracer.cpp:306-311)
    body = grab_function(gt, "GaussianTracer::createGaussianParticlesBVH")
    frag = grab_between(body, r"float\s+s\s*=\s*sqrtf", r"glm::mat4\s+transform\s*=")
    frag = frag.replace("m_gsData.particles[i]", "gp")
    open(os.path.join(out, "proxy_transform.inc"), "w").write("// GENERATED from src/GaussianTracer.cpp\n" + frag + "\n")
    # createIAS: the 12-float row-major packing (src/GaussianTracer.cpp:405-409)
    body = grab_function(gt, "GaussianTracer::createIAS")
    frag = grab_between(body, r"float\s+instance_transform\[12\]", r"\};")
    open(os.path.join(out, "ias_pack.inc"), "w").write("// GENERATED from src/GaussianTracer.cpp\n" + frag + "\n")
    # sendGeometryAttributesToDevice: the per-vertex normal transform (src/GaussianTracer.cpp:658-663)
    body = grab_function(gt, "GaussianTracer::sendGeometryAttributesToDevice")
    frag = grab_between(body, r"glm::vec3\s+glm_normal\s*=", r"vertex_normals\[i\]\s*=")
    open(os.path.join(out, "normal_transform.inc"), "w").write("// GENERATED from src/GaussianTracer.cpp\n" + frag + "\n")
    # createPlane placement weights (src/GaussianTracer.cpp:580-588)
    body = grab_function(gt, "GaussianTracer::createPlane")
    frag = grab_between(body, r"float\s+cameraWeight", r"\};")
    open(os.path.join(out, "placement.inc"), "w").write("// GENERATED from src/GaussianTracer.cpp\n" + frag + "\n")
    # defaults of initializeParams (t_min .. type), src/GaussianTracer.cpp:475-486
    body = grab_function(gt, "GaussianTracer::initializeParams")
    keep = [l for l in body.split("\n") if re.match(r"\s*params\.(t_min|t_max|minTransmittance|alpha_min|sh_degree_max|mode_fisheye|type)\s*=", l)]
    open(os.path.join(out, "param_defaults.inc"), "w").write("// GENERATED from src/GaussianTracer.cpp\n" + "\n".join(keep) + "\n")
    print("extract_ref: wrote", sorted(os.listdir(out)))


if __name__ == "__main__":
    main()
