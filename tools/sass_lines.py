"""Developer tool: join an `ncu --page source --csv --print-source sass` dump of one kernel with the line table of
`nvdisasm -g` for the same kernel, and print static instruction counts, executed instructions and stall samples per
source line (top lines) and per K1 phase (line ranges given on the command line).

usage: sass_lines.py all.sass KERNEL_MANGLED_NAME [src.csv] [file_filter]
"""
import csv, re, sys, collections

sass, kern = sys.argv[1], sys.argv[2]
src = sys.argv[3] if len(sys.argv) > 3 else None
ffilter = sys.argv[4] if len(sys.argv) > 4 else "pb_local.cuh"

# ---- address -> (file, line) innermost, and outermost line inside ffilter (inlined call chains)
addr_line = {}
infun = False
cur = None
stack_line = None
fresh = True
for ln in open(sass):
    if ln.startswith(".text."):
        infun = ln.strip() == f".text.{kern}:"
        continue
    if not infun:
        continue
    m = re.match(r'\s*//## File "([^"]+)", line (\d+)(.*)', ln)
    if m:                                   # nvdisasm -gi: one line per inline level, innermost first
        f, l, rest = m.group(1), int(m.group(2)), m.group(3)
        if fresh: cur = (f.split("/")[-1], l); stack_line = None; fresh = False
        if f.endswith(ffilter): stack_line = l
        for cf, cl in re.findall(r'inlined at "([^"]+)", line (\d+)', rest):
            if cf.endswith(ffilter): stack_line = int(cl)
        continue
    m = re.match(r'\s*/\*([0-9a-f]{4,})\*/\s+(\S.*?);', ln)
    if m and cur:
        addr_line[int(m.group(1), 16)] = (cur, stack_line, m.group(2))
        fresh = True

print(f"{len(addr_line)} instructions ({len(addr_line) * 16 / 1024:.1f} KB)")
static_inner = collections.Counter(); static_outer = collections.Counter()
for a, (c, o, t) in addr_line.items():
    static_inner[c] += 1; static_outer[o] += 1

dyn_outer = collections.Counter(); samp_outer = collections.Counter(); stall = collections.defaultdict(collections.Counter)
if src:
    rows = list(csv.reader(open(src)))
    hi = next(i for i, r in enumerate(rows) if r and r[0] == "Address")
    h = rows[hi]
    ia, ie, isamp = h.index("Address"), h.index("Instructions Executed"), h.index("# Samples")
    iwf = h.index("L1 Wavefronts Shared"); wf_outer = collections.Counter()
    stall_cols = [(i, n) for i, n in enumerate(h) if n.startswith("stall_") and "Not Issued" not in n]
    base = None
    for r in rows[hi + 1:]:
        if len(r) <= isamp: continue
        a = int(r[ia], 16) if r[ia].startswith("0x") else int(r[ia])
        if base is None: base = a
        off = a - base
        if off not in addr_line: continue
        c, o, t = addr_line[off]
        dyn_outer[o] += int(float(r[ie] or 0)); samp_outer[o] += int(float(r[isamp] or 0))
        wf_outer[o] += int(float(r[iwf] or 0))
        for i, n in stall_cols:
            v = r[i]
            if v and v != "0": stall[o][n] += int(float(v))
    tot_s = sum(samp_outer.values()) or 1; tot_d = sum(dyn_outer.values()) or 1
    print(f"total samples {tot_s}, executed warp instructions {tot_d}")
    print("line   static   dyn%   samp%   cyc/inst(rel)  top stalls")
    for o, s in sorted(samp_outer.items(), key=lambda kv: -kv[1])[:45]:
        top = ", ".join(f"{n[6:]}:{v}" for n, v in stall[o].most_common(3))
        d = dyn_outer[o]
        print(f"{str(o):>5} {static_outer[o]:7d} {100 * d / tot_d:6.2f} {100 * s / tot_s:7.2f} {(s / tot_s) / max(d / tot_d, 1e-9):8.2f}   {top}")
else:
    for o, s in static_outer.most_common(40):
        print(o, s)

# ---- per phase of local_uncut_kernel (line ranges of pb_local.cuh, override with PB_PHASES="name:lo-hi,...")
import os
ph = os.environ.get("PB_PHASES")
if not ph and len(sys.argv) > 5:          # markers "//@ph name" in the source file given as argv[5]
    names = [(i + 1, l.split("//@ph")[1].strip()) for i, l in enumerate(open(sys.argv[5])) if "//@ph" in l]
    nl = sum(1 for _ in open(sys.argv[5]))
    ph = ",".join(f"{n}:{lo}-{(names[j + 1][0] - 1) if j + 1 < len(names) else nl}" for j, (lo, n) in enumerate(names))
if src and ph:
    tot_w = sum(wf_outer.values()) or 1
    print("phase            static   dyn%   samp%   fp64-static  smem-wavefront%")
    for item in ph.split(","):
        name, rng = item.split(":"); lo, hi = map(int, rng.split("-"))
        st = sum(v for k, v in static_outer.items() if k is not None and lo <= k <= hi)
        dy = sum(v for k, v in dyn_outer.items() if k is not None and lo <= k <= hi)
        sa = sum(v for k, v in samp_outer.items() if k is not None and lo <= k <= hi)
        f64 = sum(1 for a, (c, o, t) in addr_line.items() if o is not None and lo <= o <= hi and t.split()[0].split(".")[0] in ("DFMA", "DMUL", "DADD") or
                  (o is not None and lo <= o <= hi and t.startswith("@") and len(t.split()) > 1 and t.split()[1].split(".")[0] in ("DFMA", "DMUL", "DADD")))
        wf = sum(v for k, v in wf_outer.items() if k is not None and lo <= k <= hi)
        stc = collections.Counter()
        for k_, v_ in stall.items():
            if k_ is not None and lo <= k_ <= hi: stc.update(v_)
        top = ", ".join(f"{n_[6:]}:{100 * v_ / max(sa, 1):.0f}%" for n_, v_ in stc.most_common(4))
        print(f"{name:16s} {st:6d} {100 * dy / tot_d:6.2f} {100 * sa / tot_s:7.2f} {f64:8d} {100 * wf / tot_w:10.2f}   {top}")
