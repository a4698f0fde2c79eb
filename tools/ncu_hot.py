"""Hot spots of one kernel from an .ncu-rep: the SASS rows with the most stall samples, mapped to source lines
with nvdisasm's line info.   python tools/ncu_hot.py REPORT.ncu-rep KERNEL_SUBSTRING [CUBIN_DIR] [TOP]"""
import csv, io, os, re, subprocess, sys

rep, kname = sys.argv[1], sys.argv[2]
cubdir = sys.argv[3] if len(sys.argv) > 3 else "/tmp/cubin"
top = int(sys.argv[4]) if len(sys.argv) > 4 else 40
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(out)))
hi = [i for i, r in enumerate(rows) if "# Samples" in r][0]
hdr = rows[hi]
ci = {n: hdr.index(n) for n in ("Address", "Source", "# Samples", "Instructions Executed", "stall_long_sb", "stall_barrier", "stall_wait", "stall_short_sb", "stall_branch_resolving", "stall_lg", "stall_membar", "stall_no_inst")}
data = []
for r in rows[hi + 1:]:
    if len(r) <= ci["# Samples"]:
        continue
    try:
        data.append((int(r[ci["# Samples"]] or 0), r))
    except ValueError:
        pass
# line info
lines = {}
for f in sorted(os.listdir(cubdir)):
    if not f.endswith(".cubin"):
        continue
    txt = subprocess.run(["nvdisasm", "-g", "-c", os.path.join(cubdir, f)], capture_output=True, text=True).stdout
    if kname not in txt:
        continue
    cur, infn = None, False
    for l in txt.splitlines():
        m = re.match(r"\s*\.text\.(\S+):", l)
        if m:
            infn = kname in m.group(1)
            continue
        if not infn:
            continue
        m = re.search(r'//## File "([^"]+)", line (\d+)', l)
        if m:
            cur = (os.path.basename(m.group(1)), int(m.group(2)))
            continue
        m = re.match(r"\s*/\*([0-9a-f]{4,})\*/", l)
        if m and cur:
            lines[int(m.group(1), 16)] = cur
tot = sum(d[0] for d in data)
print("total samples", tot, "instructions", sum(int(r[ci["Instructions Executed"]] or 0) for _, r in data))
base = min(int(r[ci["Address"]], 16) for _, r in data) if data else 0
byline = {}
for n, r in data:
    a = int(r[ci["Address"]], 16) - base
    byline.setdefault(lines.get(a), [0, 0])
    byline[lines.get(a)][0] += n
    byline[lines.get(a)][1] += int(r[ci["Instructions Executed"]] or 0)
print("-- by source line")
for k, (n, ie) in sorted(byline.items(), key=lambda x: -x[1][0])[:top]:
    print(f"{100 * n / max(tot, 1):5.1f}%  inst {ie:8d}  {k}")
print("-- by instruction")
for n, r in sorted(data, key=lambda x: -x[0])[:top]:
    a = int(r[ci["Address"]], 16) - base
    st = {k: r[ci[k]] for k in ("stall_long_sb", "stall_barrier", "stall_wait", "stall_short_sb", "stall_branch_resolving", "stall_membar") if r[ci[k]] not in ("0", "")}
    print(f"{100 * n / max(tot, 1):5.1f}%  {a:05x} {lines.get(a)}  exec {r[ci['Instructions Executed']]:>7}  {r[ci['Source']][:70]}  {st}")
