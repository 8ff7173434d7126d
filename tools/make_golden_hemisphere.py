"""Generates tests/golden/hemisphere20.json: the 20 hemisphere directions the reference holds as literals in
src/shaders/light_scattering.frag:133-154 (pasted there from gen_hemisphare_distrib.py:4-12 with num_pts = 20, printed as
vec3(x, z, y)) -- the one reference-held vector that pins anything about the AO pass.  The literals are parsed
from the shader where it lies; tests compare the oracle's and librc's generators with all of them, after fp32 rounding.

Run in the build container:  python tools/make_golden_hemisphere.py
"""
import json
import os
import re

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
SRC = "/root/reference/src/shaders/light_scattering.frag"


def main():
    text = open(SRC).read()
    body = text[text.index("vec3[] hemisphereDirs"):]
    body = body[:body.index("};")]
    rows = re.findall(r"vec3\(\s*([-+0-9.eE]+)\s*,\s*([-+0-9.eE]+)\s*,\s*([-+0-9.eE]+)\s*\)", body)
    assert len(rows) == 20, len(rows)
    out = {"source": "src/shaders/light_scattering.frag:133-154 (hemisphereDirs, vec3(x, z, y) as printed by "
                     "gen_hemisphare_distrib.py:11-12 with num_pts = 20)",
           "literals_x_up_y": [[a, b, c] for a, b, c in rows]}  # kept as the shader's decimal strings
    with open(os.path.join(ROOT, "tests", "golden", "hemisphere20.json"), "w") as f:
        json.dump(out, f, indent=1)
    print(len(rows), "directions")


if __name__ == "__main__":
    main()
