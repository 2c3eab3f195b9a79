"""Developer tool, run HERE after tools/capture_profiles.sh TAG: writes profiles/TAG_launches.csv,
profiles/TAG_launches_summary.txt and profiles/TAG_<kernel>_ncu_full.txt (selected raw metrics of the --set full
captures, per launch)."""
import collections, csv, os, re, shutil, subprocess, sys

tag = sys.argv[1]
root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
go = os.path.join(root, "gpurun_out")
pr = sys.argv[2] if len(sys.argv) > 2 else os.path.join(root, "profiles")
os.makedirs(pr, exist_ok=True)

# ---- launch list
src = os.path.join(go, f"{tag}_launches.csv")
if os.path.exists(src):
    if os.path.abspath(pr) != os.path.abspath(go): shutil.copy(src, os.path.join(pr, f"{tag}_launches.csv"))
    lines = [l for l in open(src) if not l.startswith("==")]
    rows = list(csv.DictReader(lines))
    agg = collections.OrderedDict()
    for r in rows:
        if r.get("Metric Name") != "gpu__time_duration.sum": continue
        name = re.sub(r"\(.*", "", r["Kernel Name"])[:100]
        v = float(r["Metric Value"].replace(",", "")); u = r["Metric Unit"]
        ms = v * {"ns": 1e-6, "us": 1e-3, "ms": 1.0, "s": 1e3}.get(u, 1e-6)
        a = agg.setdefault(name, [0, 0.0]); a[0] += 1; a[1] += ms
    tot = sum(a[1] for a in agg.values())
    with open(os.path.join(pr, f"{tag}_launches_summary.txt"), "w") as f:
        f.write(f"# {tag} launch list summary: python bench.py --steps 3 --warmup 3 --no-cpu-baseline --no-cg-tol --no-interface (ncu --metrics gpu__time_duration.sum --clock-control none, first launches of the timed program)\n")
        f.write("# kernel, launches, total ms, us/launch, share\n")
        for name, (n, ms) in sorted(agg.items(), key=lambda kv: -kv[1][1]):
            f.write(f"{name:100s} {n:5d} {ms:10.3f} ms {1e3 * ms / n:10.1f} us/launch {100 * ms / tot:5.1f}%\n")
    print(open(os.path.join(pr, f"{tag}_launches_summary.txt")).read())

# ---- full captures
KEEP = ("gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum", "gpu__dram_throughput", "launch__", "sm__issue_active", "sm__inst_executed.sum",
        "smsp__inst_executed.sum", "sm__pipe_fp64_cycles_active", "sm__warps_active", "smsp__average_warps_issue_stalled", "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum",
        "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum", "l1tex__t_sector_hit_rate", "lts__t_sector_hit_rate", "gcc__cache_requests_type_instruction.sum", "sm__icc_request",
        "lts__t_sectors_srcunit_tex_op_read.sum", "lts__t_sectors_srcunit_tex_op_write.sum", "l1tex__t_requests_pipe_lsu_mem_global_op_ld.sum", "l1tex__t_sectors_pipe_lsu_mem_global_op_ld.sum")
for fn in sorted(os.listdir(go)):
    if not (fn.startswith(tag + "_") and fn.endswith(".ncu-rep")): continue
    raw = subprocess.run(["ncu", "-i", os.path.join(go, fn), "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(raw.splitlines()))
    if len(rows) < 3: continue
    h, u, v = rows[0], rows[1], rows[2]
    out = os.path.join(pr, fn.replace(".ncu-rep", "_ncu_full.txt"))
    with open(out, "w") as f:
        f.write(f"--- {v[h.index('Kernel Name')]}   (ncu --set full --clock-control none, one launch)\n")
        for a, b, c in zip(h, u, v):
            if any(a.startswith(k) for k in KEEP): f.write(f"{a:90s} {b:15s} {c}\n")
    print("wrote", out)
