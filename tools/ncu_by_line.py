"""Development aid: join an ncu SASS source page (per-instruction counters) with nvdisasm -g line
info and print executed warp instructions + stall samples aggregated per CUDA source line.
usage: ncu_by_line.py <sass.csv from `ncu -i rep --page source --csv --print-source sass`> <nvdisasm -g output> <kernel substring>"""
import csv, re, sys, collections
sass_csv, dis, kname = sys.argv[1:4]
# --- line map from nvdisasm: per kernel section, sequence of instructions with current line
linemap = {}   # kernel section -> list of (line) per instruction index
cur = None; line = 0
for l in open(dis):
    m = re.match(r"\s*\.section\s+\.text\.(\S+?),", l)
    if m:
        cur = m.group(1); linemap[cur] = []; line = 0; continue
    m = re.search(r'//## File ".*?", line (\d+)', l)
    if m:
        line = int(m.group(1)); continue
    if cur and re.match(r"\s+/\*[0-9a-f]{4,}\*/\s+\S", l):
        linemap[cur].append(line)
# --- counters
rows = list(csv.reader(open(sass_csv)))
agg = None; out = {}
i = 0
while i < len(rows):
    r = rows[i]
    if r and r[0] == "Kernel Name":
        name = r[1]; hdr = rows[i + 1]; i += 2
        body = []
        while i < len(rows) and rows[i] and rows[i][0] != "Kernel Name":
            body.append(rows[i]); i += 1
        if kname in name and name not in out:
            out[name] = (hdr, body)
        continue
    i += 1
for name, (hdr, body) in out.items():
    sec = [k for k in linemap if kname.replace("<", "ILi").split("ILi")[0] in k]
    # pick section whose instruction count matches
    sec = [k for k in linemap if len(linemap[k]) == len(body)]
    print("==", name[:90], "instrs", len(body), "sections", len(sec))
    if not sec:
        continue
    lm = linemap[sec[0]]
    ie = hdr.index("Instructions Executed"); ss = hdr.index("# Samples")
    per = collections.defaultdict(lambda: [0, 0])
    for k, r in enumerate(body):
        per[lm[k]][0] += int(r[ie] or 0); per[lm[k]][1] += int(r[ss] or 0)
    tot_i = sum(v[0] for v in per.values()); tot_s = sum(v[1] for v in per.values())
    print("total warp instr %d samples %d" % (tot_i, tot_s))
    for ln, (a, b) in sorted(per.items(), key=lambda kv: -kv[1][0])[:int(sys.argv[4]) if len(sys.argv) > 4 else 25]:
        print("line %4d  instr %5.1f%%  samples %5.1f%%" % (ln, 100.0 * a / max(tot_i, 1), 100.0 * b / max(tot_s, 1)))
