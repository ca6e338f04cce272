"""Attribute the stall samples of an ncu report to source lines of one .cu file (SASS address -> file:line via nvdisasm -g).
Usage: python scripts/ncu_lines.py report.ncu-rep kernel-substring file.cu [top-N]"""
import collections, csv, io, os, re, subprocess, sys, tempfile
rep, kname, fname = sys.argv[1], sys.argv[2], sys.argv[3]
topn = int(sys.argv[4]) if len(sys.argv) > 4 else 40
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
lib = os.path.join(ROOT, "attackfl_b200", "lib", "libattackfl_b200.so")
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
hdr, vals = rows[0], rows[2]
want = ["gpu__time_duration.sum", "launch__registers_per_thread", "smsp__issue_active.avg.pct_of_peak_sustained_active",
        "sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active", "smsp__inst_executed.sum",
        "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_tensor.sum",
        "lts__t_sector_hit_rate.pct", "l1tex__t_sector_hit_rate.pct", "launch__grid_size", "launch__block_size",
        "sm__pipe_tensor_op_hmma_cycles_active.avg.pct_of_peak_sustained_active",
        "smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_barrier_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_short_scoreboard_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_wait_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_membar_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_lg_throttle_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_mio_throttle_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_math_pipe_throttle_per_issue_active.ratio"]
for h, u, v in zip(hdr, rows[1], vals):
    if h in want: print(f"{h} [{u}] = {v}")
tmp = tempfile.mkdtemp()
subprocess.run(["cuobjdump", "-xelf", "all", lib], cwd=tmp, capture_output=True)
sass = None
for f in os.listdir(tmp):
    if f.endswith(".cubin"):
        out = subprocess.run(["nvdisasm", "-g", "-c", os.path.join(tmp, f)], capture_output=True, text=True).stdout
        if kname in out: sass = out
cur = None; a2l = {}; a2i = {}; start = False
for ln in sass.splitlines():
    if kname in ln and ln.rstrip().endswith(":") and not ln.strip().startswith("."): start = True; a2l = {}; continue
    if not start: continue
    if ln.startswith("//-----") and a2l: break
    m = re.search(r'//## File "([^"]+)", line (\d+)(.*)', ln)
    if m:
        # innermost inlined-at chain: keep the first location that is in the wanted file
        locs = [(os.path.basename(m.group(1)), int(m.group(2)))] + [(os.path.basename(a), int(b)) for a, b in re.findall(r'inlined at "([^"]+)", line (\d+)', m.group(3))]
        cur = locs
        continue
    m = re.match(r'\s+/\*([0-9a-f]+)\*/\s+(.*);', ln)
    if m: a2l[int(m.group(1), 16)] = cur; a2i[int(m.group(1), 16)] = m.group(2).strip()
src = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(src))); hdr = rows[1]; ix = {h: i for i, h in enumerate(hdr)}; data = rows[2:]
f = lambda x: float(x) if re.fullmatch(r"[0-9.]+", x or "") else 0.0
A = lambda r: int(r[ix["Address"]], 16) if r[ix["Address"]].startswith("0x") else int(r[ix["Address"]])
base = A(data[0])
by = collections.Counter(); ins = collections.Counter(); outer = collections.Counter()
tot = 0
for r in data:
    locs = a2l.get(A(r) - base) or [("?", 0)]
    s = f(r[ix["# Samples"]]); tot += s
    inner = locs[0]
    # outermost location in the wanted file = the call site in the kernel body
    kern = next((l for l in reversed(locs) if l[0] == fname), inner)
    by[kern] += s; ins[kern] += f(r[ix["Instructions Executed"]])
print(f"total samples {tot:.0f}")
lines = open(os.path.join(ROOT, "csrc", fname)).read().splitlines()
print("--- top kernel-body lines by samples")
for (fn, l), s in by.most_common(topn):
    txt = lines[l - 1].strip()[:110] if fn == fname and 0 < l <= len(lines) else ""
    print(f"{100*s/tot:5.1f}%  {fn}:{l:<5d} inst {ins[(fn,l)]:9.0f}  {txt}")
# cumulative by 25-line buckets
print("--- by line (all, file order)")
acc = collections.OrderedDict()
for (fn, l), s in sorted(by.items(), key=lambda kv: (kv[0][0] != fname, kv[0][1])):
    if fn != fname: continue
    acc[l] = s
cum = 0
for l, s in acc.items():
    if s / tot > 0.004: print(f"  {l:5d} {100*s/tot:5.1f}%  {lines[l-1].strip()[:100]}")
