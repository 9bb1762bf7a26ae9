"""Join an ncu SASS-level source page with nvdisasm line info (same cubin) and aggregate by source function / line.

  cuobjdump -xelf all csrc/build/b2p_kernel_f32.o && nvdisasm -g -c *.cubin > f32.dis
  python tools/ncu_lines.py gpurun_out/step.ncu-rep f32.dis [--top 40]

Prints, per function of csrc/*.h (by line ranges) and per hottest source lines: executed warp instructions, stall samples,
the dominant stall reasons, and local-memory instructions (LDL/STL).
"""
import collections, csv, io, os, re, subprocess, sys
sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
from ncu_summary import functions

rep, dis = sys.argv[1], sys.argv[2]
top = int(sys.argv[sys.argv.index("--top") + 1]) if "--top" in sys.argv else 40
out = subprocess.run(["ncu", "-i", rep, "--csv", "--page", "source", "--print-source", "sass"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(out)))
h = next(r for r in rows if r and r[0] == "Address")
data = [r for r in rows if len(r) == len(h) and r[0].startswith("0x")]
# nvdisasm: ordered instructions with the preceding line marker
cur, lines = ("?", 0), []
for l in open(dis):
    m = re.search(r'//## File "([^"]+)", line (\d+)', l)
    if m:
        cur = (m.group(1).split("/")[-1], int(m.group(2))); continue
    if re.match(r"^\s+/\*[0-9a-f]+\*/", l):
        lines.append((cur, l))
# keep only the step kernel's instructions: match by count from the start of the function that contains them
if len(lines) != len(data):
    # several functions in the cubin (fma peak kernels etc.): find the contiguous run whose opcodes match
    ops = [re.sub(r"\s+", " ", r[1].strip()).split(" ")[0].lstrip("@!P0123456789 ") for r in data]
    def opc(l):
        t = re.sub(r"/\*[0-9a-f]+\*/", "", l[1]).strip().split(";")[0].split()
        t = [x for x in t if not x.startswith("@")]
        return t[0] if t else ""
    dops = [opc(l) for l in lines]
    first = data[0][1].split()[0]
    for s in range(0, len(lines) - len(data) + 1):
        if dops[s] == ops[0] and dops[s + 1] == ops[1] and dops[s + len(data) - 1] == ops[-1]:
            lines = lines[s:s + len(data)]; break
assert len(lines) == len(data), (len(lines), len(data))
ix = {k: h.index(k) for k in ("Instructions Executed", "# Samples")}
stalls = [k for k in h if k.startswith("stall_") and "Not Issued" not in k]
funcs = functions()
def fn(f, line):
    name = f
    for i, n in funcs.get(f, []):
        if i <= line: name = n
        else: break
    return name
A = collections.defaultdict(lambda: collections.Counter())
for (loc, _), r in zip(lines, data):
    for key in (("F", fn(*loc)), ("L", loc)):
        a = A[key]
        a["inst"] += int(r[ix["Instructions Executed"]] or 0); a["samp"] += int(r[ix["# Samples"]] or 0)
        if re.search(r"\b(LDL|STL)\b", r[1]): a["local"] += int(r[ix["Instructions Executed"]] or 0)
        for s in stalls: a[s] += int(r[h.index(s)] or 0)
ti = sum(a["inst"] for k, a in A.items() if k[0] == "F") or 1; ts = sum(a["samp"] for k, a in A.items() if k[0] == "F") or 1
def show(kind, n):
    for k, a in sorted(((k, a) for k, a in A.items() if k[0] == kind), key=lambda x: -x[1]["samp"])[:n]:
        st = sorted(((a[s], s[6:]) for s in stalls), reverse=True)[:3]
        print("  %-34s inst %5.1f%%  samples %5.1f%%  local %4.1f%%  | %s" % (str(k[1])[:34], 100 * a["inst"] / ti, 100 * a["samp"] / ts, 100 * a["local"] / max(1, a["inst"]),
              "  ".join("%s %.0f%%" % (n_, 100 * v / max(1, a["samp"])) for v, n_ in st)))
print("warp instructions %d, samples %d" % (ti, ts)); print("by function:"); show("F", top); print("by line:"); show("L", top)
