"""Offline attribution of an ncu report of train_kernel to source regions (SASS address -> file:line via nvdisasm -g).
Usage: python scripts/ncu_attrib.py gpurun_out/prof.ncu-rep [kernel-substring]"""
import collections, csv, io, os, re, subprocess, sys, tempfile
rep = sys.argv[1]; kname = sys.argv[2] if len(sys.argv) > 2 else "train_kernel"
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
lib = os.path.join(ROOT, "attackfl_b200", "lib", "libattackfl_b200.so")
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
hdr, vals = rows[0], rows[2]
want = ["gpu__time_duration.sum", "launch__registers_per_thread", "sm__warps_active.avg.pct_of_peak_sustained_active",
        "smsp__issue_active.avg.pct_of_peak_sustained_active", "sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active",
        "smsp__inst_executed.sum", "dram__bytes_read.sum", "dram__bytes_write.sum", "lts__t_sector_hit_rate.pct",
        "l1tex__t_sector_hit_rate.pct", "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active",
        "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "launch__grid_size", "launch__block_size",
        "smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_barrier_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_short_scoreboard_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_wait_per_issue_active.ratio",
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
cur = None; a2l = {}; start = False
for ln in sass.splitlines():
    if kname in ln and ln.rstrip().endswith(":") and not ln.strip().startswith("."): start = True; a2l = {}; continue
    if not start: continue
    if ln.startswith("//-----") and a2l: break
    m = re.search(r'//## File "([^"]+)", line (\d+)', ln)
    if m: cur = (os.path.basename(m.group(1)), int(m.group(2))); continue
    m = re.match(r'\s+/\*([0-9a-f]+)\*/\s+(.*);', ln)
    if m: a2l[int(m.group(1), 16)] = cur
src = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(src))); hdr = rows[1]; ix = {h: i for i, h in enumerate(hdr)}; data = rows[2:]
f = lambda x: float(x) if re.fullmatch(r"[0-9.]+", x or "") else 0.0
A = lambda r: int(r[ix["Address"]], 16) if r[ix["Address"]].startswith("0x") else int(r[ix["Address"]])
base = A(data[0])
fn_lines = []
for fname in ("train_core.h",):
    for i, ln in enumerate(open(os.path.join(ROOT, "csrc", fname)), 1):
        m = re.match(r"AFL_DEV \S+ (afl_\w+)\(", ln)
        if m: fn_lines.append((i, m.group(1)))
def region(k):
    if k is None: return "?"
    fn, l = k
    if fn != "train_core.h": return fn
    name = "core:top"
    for i, n in fn_lines:
        if i <= l: name = n
    return name
bi = collections.Counter(); bs = collections.Counter()
for r in data:
    k = region(a2l.get(A(r) - base)); bi[k] += f(r[ix["Instructions Executed"]]); bs[k] += f(r[ix["# Samples"]])
ti, ts = sum(bi.values()), sum(bs.values())
print(f"{'region':28s} samples%  inst%")
for k, c in bs.most_common(24): print(f"{k:28s} {c/ts*100:7.2f} {bi[k]/ti*100:7.2f}")
