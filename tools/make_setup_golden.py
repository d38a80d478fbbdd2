"""Extract the reference-held constants of the set-up into tests/golden/reference_tables.json.

Runs in the build container only (needs /root/reference).  These are literals the reference source itself
holds -- the wrf_l32 hyai / hybi tables (hybrid_coord_test_mod.F90:565-647) and the 'earth' constants
(const_mod.F90:56-66) -- parsed from the Fortran text, NOT retyped: the one part of the path where a
reference-produced datum exists.  The literals are default-real (no kind suffix), i.e. the Fortran compiler
rounds them to binary32 before widening to r8; both renderings are stored.
"""
import json, os, re, sys
import numpy as np
REF = "/root/reference/src"
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

src = open(os.path.join(REF, "dynamics/vert_coord/hybrid_coord_test_mod.F90")).read()
body = src[src.index("subroutine hybrid_coord_wrf_l32"):src.index("end subroutine hybrid_coord_wrf_l32")]
def table(name):
    blk = body[body.index(name + " = ["):]
    blk = blk[:blk.index("]")]
    vals = re.findall(r"^\s*(?:\w+ = \[)?\s*([0-9.]+)\s*(?:,\s*)?&", blk, flags=re.M)
    return vals
hyai, hybi = table("hyai"), table("hybi")
assert len(hyai) == 33 and len(hybi) == 33, (len(hyai), len(hybi))

const = open(os.path.join(REF, "utils/const_mod.F90")).read()
earth = const[const.index("case ('earth')"):const.index("case ('mars')")]
consts = {m.group(1): float(m.group(2).replace("d", "e")) for m in re.finditer(r"(\w+)\s*=\s*([0-9.]+d[-0-9]+)", earth)}
consts["omega_expr"] = "2 * pi / 86400.0d0"

out = {"generator": "tools/make_setup_golden.py from /root/reference (GMCORE @ 73c59ed)",
       "source": {"hyai_hybi": "src/dynamics/vert_coord/hybrid_coord_test_mod.F90:565-647", "const": "src/utils/const_mod.F90:56-66"},
       "wrf_l32": {"hyai_text": hyai, "hybi_text": hybi,
                   "hyai_r4": [float(np.float32(x)) for x in hyai], "hybi_r4": [float(np.float32(x)) for x in hybi]},
       "earth": consts}
json.dump(out, open(os.path.join(ROOT, "tests", "golden", "reference_tables.json"), "w"), indent=1)
print("hyai[1..3]", hyai[:3], "hybi[14]", hybi[13], consts)
