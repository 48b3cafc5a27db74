#!/usr/bin/env python3
"""Build-time extraction of the reference's hot-path routines (TEST INFRASTRUCTURE ONLY).

Cuts line ranges out of /root/reference (never committed: output goes to oracle/_ref/,
which is git-ignored) so that oracle/ref_shim.cpp can compile the reference's own text as
host C++.  Exactly two one-token shims are applied, both arithmetic-neutral (SURVEY.md
section 6.2 / Appendix A Q8):
  * kernel.cu:127  `Intersect& intersectRays` -> `Intersect intersectRays` (returned a local)
  * kernel.cu:138,182  `Ray& objectSpaceRay = Ray(...)` -> `Ray objectSpaceRay = Ray(...)`
"""
import os, re, sys

REF = os.environ.get("FIREFLY_REFERENCE", "/root/reference")
OUT = os.path.join(os.path.dirname(os.path.abspath(__file__)), "_ref")
K = "PathTracer/FireflyEngine/kernel.cu"
U = "PathTracer/FireflyEngine/utilities.h"

# (output name, file, first line, last line)  -- 1-based inclusive
RANGES = [
    ("x_sampling.inc",   U, 41, 55),    # isZero, CosineSampleHemisphere
    ("x_intersect.inc",  U, 57, 66),    # struct Intersect
    ("x_geomtypes.inc",  U, 141, 234),  # GeometryType, Triangle, Geometry
    ("x_ray.inc",        U, 257, 267),  # struct Ray
    ("x_camera.inc",     U, 269, 427),  # struct Camera
    ("x_hotpath.inc",    K, 8, 176),    # intersectPlane, intersectTriangle, setIntersection, intersectRays
    ("x_shade.inc",      K, 178, 184),  # shade (as coded: abs(normal))
    ("x_raygen.inc",     K, 198, 205),  # camera ray generation body of generateRays
]

def main():
    if not os.path.isdir(REF):
        print("reference tree not present; nothing extracted", file=sys.stderr)
        return 1
    os.makedirs(OUT, exist_ok=True)
    for name, rel, a, b in RANGES:
        with open(os.path.join(REF, rel), "r", errors="replace") as f:
            lines = f.read().splitlines()
        text = "\n".join(lines[a - 1:b]) + "\n"
        text = re.sub(r"Intersect&\s+intersectRays", "Intersect intersectRays", text)
        text = re.sub(r"Ray&\s+objectSpaceRay\s*=", "Ray objectSpaceRay =", text)
        with open(os.path.join(OUT, name), "w") as f:
            f.write("// extracted at build time from %s:%d-%d -- do not commit\n" % (rel, a, b))
            f.write(text)
    return 0

if __name__ == "__main__":
    sys.exit(main())
