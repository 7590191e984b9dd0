"""Instructions and stall samples of one kernel per barrier-delimited phase, from an ncu --set full report
(ncu -i REPORT --page source --csv --print-source sass).  usage: python tools/ncu_phases.py gpurun_out/full_x.ncu-rep"""
import csv, subprocess, sys, io

rep = sys.argv[1]
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "sass"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(out)))
hdr = rows[1]; body = rows[2:]
ia = hdr.index("Instructions Executed"); isrc = hdr.index("Source"); ism = hdr.index("# Samples")
stalls = [i for i, h in enumerate(hdr) if h.startswith("stall_") and "Not Issued" not in h]
seg = 0; tot = {}; smp = {}; ops = {}; st = {}
for r in body:
    try:
        n = int(r[ia]); s = int(r[ism])
    except ValueError:
        continue
    tot[seg] = tot.get(seg, 0) + n; smp[seg] = smp.get(seg, 0) + s
    t = r[isrc].split()
    op = (t[1] if t[0].startswith('@') else t[0]).split('.')[0]
    ops.setdefault(seg, {}); ops[seg][op] = ops[seg].get(op, 0) + n
    st.setdefault(seg, {})
    for i in stalls:
        try: st[seg][hdr[i]] = st[seg].get(hdr[i], 0) + int(r[i])
        except ValueError: pass
    if 'BAR.SYNC' in r[isrc]: seg += 1
T = sum(tot.values()); S = sum(smp.values())
for k in tot:
    top = sorted(ops[k].items(), key=lambda x: -x[1])[:8]
    tst = sorted(st[k].items(), key=lambda x: -x[1])[:5]
    print("phase %d: %9d instr %5.1f%%  samples %5.1f%%  %s" % (k, tot[k], 100 * tot[k] / T, 100 * smp[k] / max(S, 1), top))
    print("          stalls:", ["%s %.0f%%" % (a.replace("stall_", ""), 100 * b / max(smp[k], 1)) for a, b in tst])
print("total", T)
