"""Extract the reference's own known-answer tables into tests/golden/reference_golden.json.

Run in the build container (where /root/reference exists):  python tests/golden/make_reference_golden.py
Sources: apps/obstacle/results/convergence.txt (obstacle error, N x {k=0,k=1}) and apps/cuthho/cuthho.xlsx
sheet 1 (energy-norm error of cuthho_square: blocks "F.D." rows 7-10 and "Interface" rows 23-26, columns
B-D / G-I / L-N = k 0/1/2 x N 10/20/40, rows = interface refinement r 4..7).
"""
import json, os, re, zipfile
REF = "/root/reference"
out = {"source_commit": "f35d1ab", "obstacle": {}, "cuthho_fictdom": {}, "cuthho_interface": {}}
for line in open(f"{REF}/apps/obstacle/results/convergence.txt"):
    t = line.split()
    if len(t) >= 3:
        out["obstacle"][t[0]] = {"k0": float(t[1]), "k1": float(t[2])}
z = zipfile.ZipFile(f"{REF}/apps/cuthho/cuthho.xlsx")
xml = z.read("xl/worksheets/sheet1.xml").decode()
cells = {}
for m in re.finditer(r'<c r="([A-Z]+)(\d+)"([^>]*?)(?:/>|>(.*?)</c>)', xml, re.S):
    col, row, attrs, body = m.group(1), int(m.group(2)), m.group(3), m.group(4) or ""
    v = re.search(r"<v>(.*?)</v>", body)
    if v and 't="s"' not in attrs and "<f>" not in body and "<f " not in body:
        cells[(col, row)] = float(v.group(1))
cols = {0: "BCD", 1: "GHI", 2: "LMN"}
for name, r0 in (("cuthho_fictdom", 7), ("cuthho_interface", 23)):
    for k, cs in cols.items():
        for ci, N in zip(cs, (10, 20, 40)):
            for r in range(4):
                if (ci, r0 + r) in cells:
                    out[name][f"k{k}_N{N}_r{4 + r}"] = cells[(ci, r0 + r)]
path = os.path.join(os.path.dirname(os.path.abspath(__file__)), "reference_golden.json")
json.dump(out, open(path, "w"), indent=1, sort_keys=True)
print({k: (len(v) if isinstance(v, dict) else v) for k, v in out.items()})
