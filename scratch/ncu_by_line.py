"""Joins an ncu source-page CSV (SASS rows) with nvdisasm -g line info and aggregates per source line.

    cuobjdump -xelf all libsrt_b200.so ; nvdisasm -g -c srt_kernels.sm_100a.cubin > dis.txt
    ncu -i prof.ncu-rep --page source --csv > src.csv
    python scratch/ncu_by_line.py src.csv dis.txt <mangled-kernel-substring> [top]
"""
import csv
import re
import sys
from collections import defaultdict

src_csv, dis_txt, kern = sys.argv[1], sys.argv[2], sys.argv[3]
top = int(sys.argv[4]) if len(sys.argv) > 4 else 40
anchor = sys.argv[5] if len(sys.argv) > 5 else None  # attribute to the outermost frame in this file (needs nvdisasm -gi)

# offset -> (file, line) from nvdisasm
off2line = {}
cur = ("?", 0)
inside = False
in_chain = False
chain = []
for ln in open(dis_txt):
    if ln.startswith("//---") and ".text." in ln:
        inside = kern in ln
        continue
    if not inside:
        continue
    m = re.search(r'//## File "([^"]+)", line (\d+)', ln)
    if m:
        fr = (m.group(1).split("/")[-1], int(m.group(2)))
        if not in_chain:
            chain = []
            in_chain = True
        chain.append(fr)
        cur = chain[0]
        if anchor:
            for c in chain:
                if c[0] == anchor:
                    cur = c
        continue
    in_chain = False
    m = re.match(r"\s+/\*([0-9a-f]+)\*/\s+(\S.*);", ln)
    if m:
        off2line[int(m.group(1), 16)] = (cur, m.group(2).strip())

rows = list(csv.reader(open(src_csv)))
hdr = rows[1]
ix = {h: i for i, h in enumerate(hdr)}
base = None
agg = defaultdict(lambda: [0, 0, 0, 0])  # samples, inst, thread_inst, not-issued samples
tot = [0, 0, 0, 0]
stall_cols = [h for h in hdr if h.startswith("stall_") and "Not Issued" not in h]
stall_tot = defaultdict(int)
stall_by_line = defaultdict(lambda: defaultdict(int))
for r in rows[2:]:
    if len(r) < 8 or not r[0].startswith("0x"):
        continue
    a = int(r[0], 16)
    if base is None:
        base = a
    key, _ = off2line.get(a - base, (("?", 0), ""))
    v = [int(r[ix["# Samples"]] or 0), int(r[ix["Instructions Executed"]] or 0),
         int(r[ix["Thread Instructions Executed"]] or 0), int(r[ix["Warp Stall Sampling (Not-issued Samples)"]] or 0)]
    for i in range(4):
        agg[key][i] += v[i]
        tot[i] += v[i]
    for s in stall_cols:
        c = int(r[ix[s]] or 0)
        stall_tot[s] += c
        stall_by_line[key][s] += c

print("total samples %d  warp-inst %d  thread-inst %d  avg threads %.2f" % (tot[0], tot[1], tot[2], tot[2] / max(tot[1], 1)))
print("stalls:", ", ".join("%s %.1f%%" % (k[6:], 100.0 * v / max(tot[0], 1)) for k, v in sorted(stall_tot.items(), key=lambda kv: -kv[1])[:8]))
byfile = defaultdict(lambda: [0, 0, 0, 0])
for (f, l), v in agg.items():
    for i in range(4):
        byfile[f][i] += v[i]
print("\nper file:  samples%  inst%  avg-threads")
for f, v in sorted(byfile.items(), key=lambda kv: -kv[1][0]):
    print("  %-28s %5.1f%% %5.1f%%  %5.2f" % (f, 100.0 * v[0] / tot[0], 100.0 * v[1] / tot[1], v[2] / max(v[1], 1)))
print("\ntop lines: samples%  inst%  avg-threads  top stalls")
for (f, l), v in sorted(agg.items(), key=lambda kv: -kv[1][0])[:top]:
    st = sorted(stall_by_line[(f, l)].items(), key=lambda kv: -kv[1])[:3]
    print("  %-24s:%4d %5.1f%% %5.1f%%  %5.2f  %s" % (f, l, 100.0 * v[0] / tot[0], 100.0 * v[1] / tot[1], v[2] / max(v[1], 1),
                                                   " ".join("%s=%d" % (k[6:], c) for k, c in st)))
