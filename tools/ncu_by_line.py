"""Join an ncu source-page CSV (per SASS address) with nvdisasm --print-line-info output (per address -> source line).
usage: ncu_by_line.py src.csv disasm.txt [min_pct]"""
import csv, re, sys, collections
src, dis = sys.argv[1], sys.argv[2]
minpct = float(sys.argv[3]) if len(sys.argv) > 3 else 0.5
line_of = {}
cur = None
for l in open(dis):
    m = re.search(r'//## File "([^"]+)", line (\d+)', l)
    if m:
        # inlined-at chains: keep the first (innermost) only if from our file; prefer outermost our-file line
        cur = (m.group(1).split("/")[-1], int(m.group(2)))
        continue
    m = re.match(r'\s+/\*([0-9a-f]{4,6})\*/', l)
    if m:
        line_of[int(m.group(1), 16)] = cur
rows = list(csv.reader(open(src)))
hdr = rows[1]; data = rows[2:]
for _k, _r in enumerate(data):                      # several launches in one export: keep the first
    if _r and _r[0] == "Kernel Name":
        data = data[:_k]
        break
ia = hdr.index("Address"); ie = hdr.index("Instructions Executed"); isamp = hdr.index("# Samples")
stalls = [h for h in hdr if h.startswith("stall_") and "Not Issued" not in h]
sidx = [hdr.index(h) for h in stalls]
base = int(data[0][ia], 16)
agg = collections.defaultdict(lambda: [0, 0] + [0] * len(stalls))
tot = ts = 0
for r in data:
    off = int(r[ia], 16) - base
    k = line_of.get(off, ("?", 0))
    a = agg[k]
    n, s = int(r[ie]), int(r[isamp])
    a[0] += n; a[1] += s; tot += n; ts += s
    for j, i in enumerate(sidx):
        a[2 + j] += int(r[i])
print("total warp instructions", tot, "samples", ts)
srcs = {}
for (f, ln), a in sorted(agg.items()):
    if a[0] >= tot * minpct / 100 or a[1] >= ts * minpct / 100:
        top = sorted(((a[2 + j], stalls[j][6:]) for j in range(len(stalls))), reverse=True)[:3]
        tops = " ".join(f"{n}:{100 * c / ts:.1f}" for c, n in top if c)
        text = ""
        try:
            if f not in srcs:
                import glob
                p = glob.glob(f"/root/repo/k8s_rocm_metrics_dashboard_b200/csrc/{f}")
                srcs[f] = open(p[0]).read().split("\n") if p else []
            text = srcs[f][ln - 1].strip()[:90] if srcs[f] else ""
        except Exception:
            pass
        print(f"{f[:22]:22s}:{ln:4d} inst {100 * a[0] / tot:5.1f}% samp {100 * a[1] / ts:5.1f}%  [{tops}]  {text}")
