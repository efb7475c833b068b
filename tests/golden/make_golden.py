#!/usr/bin/env python
"""Extracts the reference's own golden vectors for the hot path into tests/golden/*.json.

Run in the build container, where /root/reference exists (the GPU box has no reference
tree; tests read only the committed JSON):

    python tests/golden/make_golden.py

Sources:
  examples/tmp.txt:15-33                     Jacobian MatView, 16x16, 2 ranks
  examples/SNES_test/elastic/*_output_*.vtk  nodal stress + displacement, elastic
  examples/CZM_for_instability_test.log      UEL per-ip cohesive nodal forces
  examples/{lin_two_quads_qs,cohesive_test_2,two_quads_qs}.inp  input decks (data)
"""
import json
import os
import re

REF = "/root/reference/examples"
OUT = os.path.dirname(os.path.abspath(__file__))


def matview(path):
    rows = {}
    started = False
    for line in open(path):
        if line.startswith("Mat Object"):
            if started:
                break
            started = True
            continue
        m = re.match(r"row (\d+):(.*)", line)
        if started and m:
            ents = re.findall(r"\((\d+), ([^)]+)\)", m.group(2))
            rows[int(m.group(1))] = [[int(c), float(v)] for c, v in ents]
    return [rows[i] for i in range(len(rows))]


def vtk_fields(path):
    lines = open(path).read().splitlines()
    out = {}
    i = 0
    while i < len(lines):
        if lines[i].startswith("POINTS"):
            n = int(lines[i].split()[1])
            out["points"] = [[float(x) for x in lines[i + 1 + k].split()] for k in range(n)]
        if lines[i].startswith("SCALARS STRESS"):
            out["stress"] = [[float(x) for x in lines[i + 2 + k].split()] for k in range(n)]
        if lines[i].startswith("VECTORS displacements"):
            out["disp"] = [[float(x) for x in lines[i + 1 + k].split()] for k in range(n)]
        i += 1
    return out


def uel_forces(path):
    """All 'Cohesive Forces are' blocks as [8] vectors (node-major, (x, y) per node)."""
    lines = open(path, errors="replace").read().splitlines()
    blocks = []
    for i, l in enumerate(lines):
        if "Cohesive Forces" in l:
            vals = []
            for k in range(8):
                toks = lines[i + 1 + k].split()
                vals.append(float(toks[2].replace("E+", "E+").replace("E-", "E-")))
            blocks.append(vals)
    nz = [b for b in blocks if any(abs(v) > 0 for v in b)]
    return nz[:8]


def main():
    gold = {
        "source": "srinath-chakravarthy/FEM examples/ (see make_golden.py docstring)",
        "jacobian_matview_cohesive_test_2": matview(f"{REF}/tmp.txt"),
        "elastic_vtk_rank0": vtk_fields(f"{REF}/SNES_test/elastic/0_output_000000.vtk"),
        "elastic_vtk_rank1": vtk_fields(f"{REF}/SNES_test/elastic/1_output_000000.vtk"),
        # raw text of the same two files: the byte-for-byte target of fem_b200.mesh.write_vtk (WriteOutput, src/m_io.F90:480-555)
        "elastic_vtk_rank0_text": open(f"{REF}/SNES_test/elastic/0_output_000000.vtk").read(),
        "elastic_vtk_rank1_text": open(f"{REF}/SNES_test/elastic/1_output_000000.vtk").read(),
        "uel_ip_forces": uel_forces(f"{REF}/CZM_for_instability_test.log"),
        "decks": {n: open(f"{REF}/{n}.inp").read()
                  for n in ("lin_two_quads_qs", "cohesive_test_2", "two_quads_qs")},
    }
    with open(os.path.join(OUT, "reference_golden.json"), "w") as fh:
        json.dump(gold, fh, indent=1)
    print("wrote", os.path.join(OUT, "reference_golden.json"))


if __name__ == "__main__":
    main()
