#!/usr/bin/env python
"""Extracts the solve-level golden vectors of the reference into JSON fixtures.

Run in the build container (the reference tree is not on the GPU box):
    python tests/golden/make_golden.py [/root/reference]
Each fixture holds the *input data* of one reference test (nodes, connectivity, material, surface
conditions, time stepping) and its RESULT DESCRIPTION (node, quantity, value, tolerance).  No reference
code is copied; tests/structure_driver.py replays the problem on the oracle and on the CUDA library.
"""
import json
import os
import re
import sys

import yaml

REF = sys.argv[1] if len(sys.argv) > 1 else "/root/reference"
INP = os.path.join(REF, "tests", "input_files")
OUT = os.path.dirname(os.path.abspath(__file__))

CASES = [
    "solid_ele_hex8_Standard_eas_none_stressout.4C.yaml",
    "solid_ele_hex8_Standard_eas_none_dynamic.dat",
    "solid_ele_hex27_Standard_stressout.dat",
    "solid_ele_hex27_Standard_dynamic.dat",
    "solid_ele_tet10_Standard_stressout.dat",
    "solid_ele_tet10_Standard_dynamic.dat",
    "sohex8fbar_cooks_nl_new_struc.dat",  # 64 hex8 TECH fbar, CoupNeoHooke NUE 0.4999, statics, 20 load steps
    "solid_ele_hex8_Standard_eas_none_volume_neumann.dat",  # surface + VOLUME Neumann loads with functions of space and time
    "solid_ele_hex8_Standard_fbar.dat",  # TECH fbar, MAT_ElastHyper with TWO summands: CoupNeoHooke + CoupAnisoExpo (FIBER1 on the element)
    # SURVEY 8f F4 cell types (generic kernel): GenAlpha with surface loads on quad8 / tri3 / quad4 faces
    "solid_ele_hex20_Standard_dynamic.dat",
    "solid_ele_tet4_Standard_dynamic.dat",
    "solid_ele_wedge6_Standard_dynamic.dat",
    "solid_ele_pyramid5_Standard_dynamic.dat",
    "solid_ele_pyramid5_Standard_fbar.dat",  # F-bar on pyramid5 + CoupAnisoExpo fibres, statics, surface + volume loads
    "solid_ele_hex20_Standard_volume_neumann.dat",
    "solid_ele_tet4_Standard_volume_neumann.dat",
    "solid_ele_pyramid5_Standard_volume_neumann.dat",
    # nodal Cauchy stresses of the F4 cells (Gauss-point extrapolation per base cell; pyramid5: the reference's fixed matrix)
    "solid_ele_hex20_Standard_stressout.dat",
    "solid_ele_tet4_Standard_stressout.dat",
    "solid_ele_wedge6_Standard_stressout.dat",
    "solid_ele_pyramid5_Standard_stressout.dat",
    # MAT_ElastHyper = CoupNeoHooke + CoupAnisoExpo with FIBER1 per element, displacement-based, all seven cells
    "solid_ele_hex8_Standard_eas_none_element_fibers.dat",
    "solid_ele_hex27_Standard_element_fibers.dat",
    "solid_ele_tet10_Standard_element_fibers.dat",
    "solid_ele_hex20_Standard_element_fibers.dat",
    "solid_ele_tet4_Standard_element_fibers.dat",
    "solid_ele_wedge6_Standard_element_fibers.dat",
    "solid_ele_pyramid5_Standard_element_fibers.dat",
    # KINEM linear (DisplacementBasedLinearKinematicsFormulation) with St.Venant-Kirchhoff, surface + volume loads
    "solid_ele_hex8_Standard_linear.dat",
    "solid_ele_hex20_Standard_linear.dat",
    "solid_ele_tet4_Standard_linear.dat",
    "solid_ele_pyramid5_Standard_linear.dat",
    # CoupAnisoExpo with INIT 3: FIBER1 at the nodes (FNODE), interpolated to the Gauss points
    "solid_ele_hex8_Standard_eas_none_nodal_fibers.dat",
    "solid_ele_tet4_Standard_nodal_fibers.dat",
    # strain output (Green-Lagrange, Euler-Almansi, logarithmic) next to 2nd Piola-Kirchhoff / Cauchy stresses
    "solid_ele_hex8_Standard_stressout_2pk_strainout_gl.dat",
    "solid_ele_hex8_Standard_stressout_cauchy_strainout_ea.dat",
    "solid_ele_hex8_Standard_stressout_cauchy_strainout_log.dat",
    # TECH eas_mild / eas_full (9 / 21 enhanced strain modes, condensed on the element): displacements, nodal stresses and strains
    "solid_ele_hex8_Standard_eas_mild_stressout_2pk_strainout_gl.dat",
    "solid_ele_hex8_Standard_eas_mild_stressout_cauchy_strainout_ea.dat",
    "solid_ele_hex8_Standard_eas_full_stressout_2pk_strainout_gl.dat",
    "solid_ele_hex8_Standard_eas_full_stressout_cauchy_strainout_ea.dat",
    "solid_ele_tet10_Standard_nodal_fibers.dat",
    # PRESTRESS mulf (PRESTRESS_TECH mulf): two prestress steps, then eight load steps; displacements and nodal stresses
    "solid_ele_hex8_Standard_mulf.dat",
    "solid_ele_hex20_Standard_mulf.dat",
    "solid_ele_tet4_Standard_mulf.dat",
    "solid_ele_wedge6_Standard_mulf.dat",
    "solid_ele_pyramid5_Standard_mulf.dat",
    "solid_ele_hex8_Standard_mulf_fbar.dat",  # PRESTRESS_TECH mulf with TECH fbar
    "solid_ele_hex27_Standard_mulf.dat",
    "solid_ele_tet10_Standard_mulf.dat",
    # more inputs of the same families
    "solid_ele_hex27_Standard_linear.dat",
    "solid_ele_hex27_Standard_volume_neumann.dat",
    "solid_ele_tet10_Standard_volume_neumann.dat",
    "solid_ele_tet4_Standard_stressout_2pk_strainout_gl.dat",
    "solid_ele_tet4_Standard_stressout_cauchy_strainout_ea.dat",
    "solid_ele_tet4_Standard_stressout_cauchy_strainout_log.dat",
    # hex18 (quad9 x line2, hex_18point; quad6 side faces)
    "solid_ele_hex18_Standard_dynamic.dat",
    "solid_ele_hex18_Standard_element_fibers.dat",
    "solid_ele_hex18_Standard_volume_neumann.dat",
    "solid_ele_hex8_Standard_shell_eas.dat",  # TECH shell_eas = EAS with the 7 modes of eas_sosh8
    # Cook's membrane with 64 EAS elements, near-incompressible Neo-Hooke, 100 load steps (the multi-element EAS golden)
    "sohex8_easmild_cooks_nl.dat",
    # 125 eas_full elements, St.Venant-Kirchhoff nu = 0.4999, symmetry planes with LINE Dirichlet conditions that override the
    # surface ones; patch tests with a POINT condition
    # nurbs27: one element and a 2 x 3 x 4 patch (control points with weights, STRUCTURE KNOTVECTORS)
    "solid_ele_nurbs27_Standard_Static.dat",
    "solid_ele_nurbs27_Standard_Static_multielement.dat",  # no RESULT DESCRIPTION: CUDA against oracle only
    "solid_runtime_nurbs.dat",  # two curved patches (17 elements), weights != 1, three load steps
    # the runtime-output family: St.Venant-Kirchhoff E = 100, nu = 0, two load steps, every node's displacement listed
    "solid_runtime_hex8.dat",
    "solid_runtime_hex20.dat",
    "solid_runtime_hex27.dat",
    "solid_runtime_tet4.dat",
    "solid_runtime_tet10.dat",
    "solid_runtime_wedge6.dat",
    "solid_runtime_pyramid5.dat",
    # patch tests on tetrahedra
    "sotet10_patchtest.dat",
    "sotet4_patchtest.dat",
    # found by scanning the reference's input directory for Structure problems made of SOLID elements, in-scope materials and
    # conditions only: linear patch tests, single-step solver tests, body loads per cell, and the same element inputs under
    # INT_STRATEGY Old (the elements are the new ones; another time-integration framework drives them)
    "patch_test_cube_h27_linear_no_override_dbc.dat",
    "patch_test_cube_linear_no_override_dbc.dat",
    "patch_test_cube_linear_test_react.dat",
    "patch_test_cube_h27_linear_test_react.dat",
    "sohex27_patchtest_nl_cost_drt.dat",
    "structure_new_nln_solver_singlestep_cube_linear.dat",
    "structure_new_nln_solver_singlestep_cube_linear_restart.dat",
    "solid_runtime_material_element_id.dat",
    "solid_ele_nurbs27_solid_runtime.dat",
    "solidh8_ele_standard_volume_neumann.dat",
    "solidh18_ele_standard_volume_neumann.dat",
    "solidh20_ele_standard_volume_neumann.dat",
    "solidh27_ele_standard_volume_neumann.dat",
    "solidt4_ele_standard_volume_neumann.dat",
    "solidt10_ele_standard_volume_neumann.dat",
    "solid_ele_hex8_Old_eas_none_dynamic.dat",
    "solid_ele_hex8_Old_eas_none_volume_neumann.dat",
    "solid_ele_hex8_Old_eas_none_element_fibers.dat",
    "solid_ele_hex8_Old_eas_none_nodal_fibers.dat",
    "solid_ele_hex8_Old_eas_mild_dynamic.dat",
    "solid_ele_hex8_Old_eas_mild_volume_neumann.dat",
    "solid_ele_hex8_Old_eas_mild_element_fibers.dat",
    "solid_ele_hex8_Old_eas_mild_nodal_fibers.dat",
    "solid_ele_hex8_Old_eas_full_dynamic.dat",
    "solid_ele_hex8_Old_eas_full_volume_neumann.dat",
    "solid_ele_hex8_Old_eas_full_nodal_fibers.dat",
    "solid_ele_hex8_Old_fbar.dat",
    "solid_ele_hex8_Old_mulf.dat",
    "solid_ele_hex18_Old_dynamic.dat",
    "solid_ele_hex18_Old_volume_neumann.dat",
    "solid_ele_hex18_Old_element_fibers.dat",
    "solid_ele_hex20_Old_dynamic.dat",
    "solid_ele_hex20_Old_volume_neumann.dat",
    "solid_ele_hex20_Old_element_fibers.dat",
    "solid_ele_hex27_Old_dynamic.dat",
    "solid_ele_hex27_Old_volume_neumann.dat",
    "solid_ele_hex27_Old_element_fibers.dat",
    "solid_ele_pyramid5_Old_dynamic.dat",
    "solid_ele_pyramid5_Old_volume_neumann.dat",
    "solid_ele_pyramid5_Old_element_fibers.dat",
    "solid_ele_pyramid5_Old_fbar.dat",
    "solid_ele_tet10_Old_dynamic.dat",
    "solid_ele_tet10_Old_volume_neumann.dat",
    "solid_ele_tet10_Old_element_fibers.dat",
    "solid_ele_tet10_Old_nodal_fibers.dat",
    "solid_ele_tet4_Old_dynamic.dat",
    "solid_ele_tet4_Old_volume_neumann.dat",
    "solid_ele_tet4_Old_element_fibers.dat",
    "solid_ele_tet4_Old_nodal_fibers.dat",
    "solid_ele_wedge6_Old_dynamic.dat",
    "solid_ele_wedge6_Old_element_fibers.dat",
    # prescribed displacements (VAL * FUNCT with a VARIABLE of TYPE linearinterpolation), KINEM linear beams with analytical solutions
    "patch_test_cube_linear_prescribed_disp_no_override_dbc.dat",
    "patch_test_cube_linear_prescribed_disp_triaxial_no_override_dbc.dat",
    "error_analytical_beam_cantilever_end_surface_load_with_poissons_effect.dat",
    "error_analytical_beam_cantilever_with_bodyload_without_poissons_effect.dat",
    "sohex20_patchtest_nl_cost_drt.dat",  # hex20 elements on the node set of a hex27 mesh: the unused nodes are dropped by the replay
    "sosh8_cooks_nl.dat",  # Cook's membrane with 16 eas_mild elements
    "sohex8_easmild_cooks_nl_dynsta.dat",  # the 64-element one with TOLRES 1e-6
    "loadlin.dat",  # 300 eas_full hex8 under TYPE orthopressure with LOADLIN yes (follower load and its stiffness)
    "prestress_hex27.dat",  # the same on 144 hex27 (quad9 faces) and
    "prestress_tet10.dat",  # on 1029 tet10 (tri6 faces)
    "prestress_hex20.dat",  # 144 hex20 with MULF under a pseudo_orthopressure load (pressure along the last converged normal)
    "soweg6_disp_patchtest.dat",  # 4 hex8 + 6 wedge6 in one discretization (tests/test_mixed_cells.py)
    "prestress_new.dat",  # six bodies: tet4, tet10, hex8, hex8 F-bar, hex20, hex27, all with PRESTRESS_TECH mulf (tests/test_mixed_cells.py)
    "sohex8_incompr_block.dat",
    "sohex8_incompr_block_nln.dat",
    "sohex8_easfull_patchtest.dat",
    "sohex8_disp_patchtest.dat",
    "sohex8_easmild_cooks_lin.dat",  # the same mesh with KINEM linear, St.Venant-Kirchhoff nu = 0.4999, one load step
    "sohex8_easfull_cooks_lin.dat",
    # solid shells: assumed natural strains, alone and with the EAS modes of the solid shell
    "solid_ele_hex8_Standard_shell_ans.dat",
    "solid_ele_hex8_Standard_shell_eas_ans.4C.yaml",
    "solid_ele_wedge6_Standard_shell_ans.4C.yaml",
    "solid_ele_wedge6_Standard_shell_eas_ans.4C.yaml",
]


def dat_sections(path):
    sec, cur = {}, None
    for line in open(path):
        line = line.rstrip("\n")
        m = re.match(r"^-{3,}(.*)$", line)
        if m:
            cur = m.group(1).strip()
            sec[cur] = []
        elif cur is not None and line.strip() and not line.startswith("//"):
            sec[cur].append(line.strip())
    return sec


def keyvals(lines):
    out = {}
    for l in lines:
        parts = l.split()
        out[parts[0]] = " ".join(parts[1:])
    return out


def parse_condition_dat(line):
    t = line.split()
    e = int(t[t.index("E") + 1])
    n = int(t[t.index("NUMDOF") + 1])
    i = t.index("ONOFF")
    onoff = [int(x) for x in t[i + 1:i + 1 + n]]
    i = t.index("VAL")
    val = [float(x) for x in t[i + 1:i + 1 + n]]
    i = t.index("FUNCT")
    funct = [int(x) if x.lstrip("-").isdigit() else 0 for x in t[i + 1:i + 1 + n]]
    typ = t[t.index("TYPE") + 1] if "TYPE" in t else "Live"
    return {"E": e, "onoff": onoff, "val": val, "funct": funct, "type": typ}


def from_dat(path):
    s = dat_sections(path)
    dyn = keyvals(s["STRUCTURAL DYNAMIC"])
    mats = {}
    for l in s["MATERIALS"]:
        t = l.split()
        mats[int(t[1])] = t[2:]
    functs = {int(k[5:]): re.split(r"SYMBOLIC_FUNCTION_OF_(?:SPACE_)?TIME", v[0])[1].strip() for k, v in s.items() if re.match(r"FUNCT\d+$", k)}
    # VARIABLE lines of a function: piecewise linear in time (TYPE linearinterpolation NUMPOINTS n TIMES ... VALUES ...)
    funct_vars = {}
    for k, v in s.items():
        if re.match(r"FUNCT\d+$", k):
            for l in v[1:]:
                t = l.split()
                if t[0] == "VARIABLE" and t[t.index("TYPE") + 1] == "linearinterpolation":
                    n = int(t[t.index("NUMPOINTS") + 1])
                    i, j = t.index("TIMES"), t.index("VALUES")
                    funct_vars.setdefault(int(k[5:]), {})[t[t.index("NAME") + 1]] = {
                        "times": [float(x) for x in t[i + 1:i + 1 + n]], "values": [float(x) for x in t[j + 1:j + 1 + n]]}
    results = []
    for l in s.get("RESULT DESCRIPTION", []):
        t = l.split()
        results.append({"node": int(t[t.index("NODE") + 1]) if "NODE" in t else -1, "quantity": t[t.index("QUANTITY") + 1],
                        "value": float(t[t.index("VALUE") + 1]), "tolerance": float(t[t.index("TOLERANCE") + 1])})
    surf = {}
    for l in s.get("DSURF-NODE TOPOLOGY", []):
        t = l.split()
        surf.setdefault(int(t[3]), []).append(int(t[1]))
    line, point = {}, {}
    for l in s.get("DLINE-NODE TOPOLOGY", []):
        t = l.split()
        line.setdefault(int(t[3]), []).append(int(t[1]))
    for l in s.get("DNODE-NODE TOPOLOGY", []):
        t = l.split()
        point.setdefault(int(t[3]), []).append(int(t[1]))
    vol = {}
    for l in s.get("DVOL-NODE TOPOLOGY", []):
        t = l.split()
        vol.setdefault(int(t[3]), []).append(int(t[1]))
    nodes, node_fibers, weights = {}, {}, {}
    for l in s["NODE COORDS"]:
        t = l.split()
        nodes[int(t[1])] = [float(t[3]), float(t[4]), float(t[5])]
        if t[0] == "CP":  # NURBS control point: CP id COORD x y z weight
            weights[int(t[1])] = float(t[6])
        if t[0] == "FNODE":  # fibre node: FNODE id COORD x y z FIBER1 a b c
            i = t.index("FIBER1")
            node_fibers[int(t[1])] = [float(x) for x in t[i + 1:i + 4]]
    eles = []
    for l in s["STRUCTURE ELEMENTS"]:
        t = l.split()
        i = t.index("MAT")
        eles.append({"id": int(t[0]), "type": t[2], "nodes": [int(x) for x in t[3:i]], "mat": int(t[i + 1]),
                     "opts": " ".join(t[i + 2:])})
    # STRUCTURE KNOTVECTORS: patches of NURBS_DIMENSION knot vectors (NUMKNOTS, DEGREE, TYPE, then the knots)
    patches = []
    kv = s.get("STRUCTURE KNOTVECTORS", [])
    i = 0
    while i < len(kv):
        t = kv[i].split()
        if t[0] == "BEGIN":
            patches.append([])
        elif t[0] == "NUMKNOTS":
            n = int(t[1])
            deg, typ = int(kv[i + 1].split()[1]), kv[i + 2].split()[1]
            patches[-1].append({"degree": deg, "type": typ, "knots": [float(x) for x in kv[i + 3:i + 3 + n]]})
            i += 2 + n
        i += 1
    return {"dynamic": dyn, "io": keyvals(s.get("IO", [])), "materials": mats, "weights": weights, "knotvectors": patches, "funct_vars": funct_vars, "functs": functs, "results": results, "dsurf": surf,
            "dirich": [parse_condition_dat(l) for l in s.get("DESIGN SURF DIRICH CONDITIONS", [])],
            "dline": line, "dirich_line": [parse_condition_dat(l) for l in s.get("DESIGN LINE DIRICH CONDITIONS", [])],
            "dnode": point, "dirich_point": [parse_condition_dat(l) for l in s.get("DESIGN POINT DIRICH CONDITIONS", [])],
            "neumann": [parse_condition_dat(l) for l in s.get("DESIGN SURF NEUMANN CONDITIONS", [])],
            "dvol": vol, "vol_neumann": [parse_condition_dat(l) for l in s.get("DESIGN VOL NEUMANN CONDITIONS", [])],
            "nodes": nodes, "node_fibers": node_fibers, "elements": eles}


def from_yaml(path):
    y = yaml.safe_load(open(path))
    dyn = {k: str(v) for k, v in y["STRUCTURAL DYNAMIC"].items()}
    mats = {}
    for m in y["MATERIALS"]:
        mid = m["MAT"]
        name = [k for k in m if k != "MAT"][0]
        flat = [name]
        for k, v in m[name].items():
            flat += [k] + ([str(x) for x in v] if isinstance(v, list) else [str(v)])
        mats[mid] = flat
    functs = {int(k[5:]): v[0]["SYMBOLIC_FUNCTION_OF_SPACE_TIME"] for k, v in y.items() if k.startswith("FUNCT")}
    results = [{"node": r["STRUCTURE"]["NODE"], "quantity": r["STRUCTURE"]["QUANTITY"], "value": float(r["STRUCTURE"]["VALUE"]),
                "tolerance": float(r["STRUCTURE"]["TOLERANCE"])} for r in y["RESULT DESCRIPTION"]]
    surf = {}
    for l in y["DSURF-NODE TOPOLOGY"]:
        t = l.split()
        surf.setdefault(int(t[3]), []).append(int(t[1]))
    nodes = {}
    for l in y["NODE COORDS"]:
        t = l.split()
        nodes[int(t[1])] = [float(t[3]), float(t[4]), float(t[5])]
    eles = []
    for l in y["STRUCTURE ELEMENTS"]:
        t = l.split()
        i = t.index("MAT")
        eles.append({"id": int(t[0]), "type": t[2], "nodes": [int(x) for x in t[3:i]], "mat": int(t[i + 1]),
                     "opts": " ".join(t[i + 2:])})

    def cond(c):
        return {"E": c["E"], "onoff": c["ONOFF"], "val": [float(v) for v in c["VAL"]],
                "funct": [int(f) if f else 0 for f in c["FUNCT"]], "type": c.get("TYPE", "Live")}
    return {"dynamic": dyn, "materials": mats, "functs": functs, "results": results, "dsurf": surf,
            "dirich": [cond(c) for c in y.get("DESIGN SURF DIRICH CONDITIONS", [])],
            "neumann": [cond(c) for c in y.get("DESIGN SURF NEUMANN CONDITIONS", [])],
            "nodes": nodes, "elements": eles}


def main():
    for name in CASES:
        path = os.path.join(INP, name)
        d = from_yaml(path) if name.endswith(".yaml") else from_dat(path)
        d["source"] = "tests/input_files/" + name
        # keep only what the in-scope path can replay: ElastHyper with CoupNeoHooke / CoupSVK / CoupAnisoExpo summands
        # or MAT_Struct_StVenantKirchhoff
        m1 = d["materials"][1]
        assert m1[0] in ("MAT_ElastHyper", "MAT_Struct_StVenantKirchhoff"), name
        if m1[0] == "MAT_ElastHyper":
            nm = int(m1[m1.index("NUMMAT") + 1])
            for mid in m1[m1.index("MATIDS") + 1:m1.index("MATIDS") + 1 + nm]:
                assert d["materials"][int(mid)][0] in ("ELAST_CoupNeoHooke", "ELAST_CoupSVK", "ELAST_CoupAnisoExpo"), name
        out = os.path.join(OUT, re.sub(r"\.(dat|4C\.yaml)$", "", name) + ".json")
        with open(out, "w") as f:
            json.dump(d, f, indent=0, sort_keys=True)
        print(out, len(d["nodes"]), "nodes", len(d["elements"]), "elements", len(d["results"]), "results")




# ---- solid_runtime_ghosting_output.dat: reference result = runtime VTK output (binary .vtu per rank) -------------
def _vtu_arrays(path):
    """DataArrays of a vtkZLibDataCompressor / base64 .vtu (header: nblocks, blocksize, lastsize, csizes...; uint32)"""
    import base64
    import struct
    import zlib
    import xml.etree.ElementTree as ET
    import numpy as np
    out = {}
    root = ET.parse(path).getroot()
    dt = {"Float64": np.float64, "Int32": np.int32, "UInt8": np.uint8}
    for da in root.iter("DataArray"):
        if da.get("format") != "binary":
            continue
        txt = "".join(da.text.split())
        nb = struct.unpack("<I", base64.b64decode(txt[:8])[:4])[0]
        hlen = 4 * (3 + nb)
        henc = (hlen + 2) // 3 * 4
        hdr = struct.unpack("<%dI" % (3 + nb), base64.b64decode(txt[:henc])[:hlen])
        raw = base64.b64decode(txt[henc:])
        data, pos = b"", 0
        for cs in hdr[3:]:
            data += zlib.decompress(raw[pos:pos + cs])
            pos += cs
        arr = np.frombuffer(data, dtype=dt[da.get("type")])
        nc = int(da.get("NumberOfComponents", "1"))
        out[da.get("Name") or "points"] = arr.reshape(-1, nc) if nc > 1 else arr
    return out


def ghosting_case():
    import glob
    import numpy as np
    name = "solid_runtime_ghosting_output.dat"
    d = from_dat(os.path.join(INP, name))
    assert d["materials"][1][0] == "MAT_Struct_StVenantKirchhoff"
    d["source"] = "tests/input_files/" + name + " + tests/input_files/ref/solid_runtime_ghosting_output-vtk/vtk-files/structure-000002-*.vtu"
    coords = {tuple(np.round(v, 9)): k for k, v in d["nodes"].items()}
    disp = {}
    for f in sorted(glob.glob(os.path.join(INP, "ref", "solid_runtime_ghosting_output-vtk", "vtk-files", "structure-000002-*.vtu"))):
        a = _vtu_arrays(f)
        for x, u in zip(a["points"], a["displacement"]):
            nid = coords[tuple(np.round(x, 9))]
            if nid in disp:
                assert np.allclose(disp[nid], u, atol=1e-13)
            disp[nid] = [float(v) for v in u]
    assert len(disp) == len(d["nodes"])
    d["results"] = [{"node": nid, "quantity": q, "value": disp[nid][c], "tolerance": 1e-8}
                    for nid in sorted(disp) if nid % 4 == 1 for c, q in enumerate(("dispx", "dispy", "dispz"))]  # every 4th node
    out = os.path.join(OUT, "solid_runtime_ghosting_output.json")
    with open(out, "w") as f:
        json.dump(d, f, indent=0, sort_keys=True)
    print(out, len(d["nodes"]), "nodes", len(d["elements"]), "elements", len(d["results"]), "results")


if __name__ == "__main__":
    main()
    ghosting_case()
