"""Summarise `ncu --page source --print-source cuda,sass --csv` per source line: stall samples and
executed warp instructions.  usage: ncu_lines.py report.ncu-rep [top_n]"""
import csv, subprocess, sys, collections

rep = sys.argv[1]
top = int(sys.argv[2]) if len(sys.argv) > 2 else 40
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "cuda,sass"],
                     capture_output=True, text=True).stdout
rows = list(csv.reader(out.splitlines()))
cur_file = ""
hdr = None
lines = []
for r in rows:
    if not r:
        continue
    if r[0] == "File Path":
        cur_file = r[1].split("/")[-1]
    elif r[0] == "Line No":
        hdr = r
    elif hdr and r[0] not in ("", "Function Name") and r[0].isdigit():
        d = dict(zip(hdr[4:], r[4:]))
        stalls = {k: int(v) for k, v in d.items() if k.startswith("stall_") and "Not Issued" not in k and v.isdigit() and int(v)}
        num = lambda k: int(d[k]) if d.get(k, "").isdigit() else 0
        lines.append((cur_file, int(r[0]), r[1].strip(), num("# Samples"), num("Instructions Executed"), stalls))
tot_s = sum(l[3] for l in lines) or 1
tot_i = sum(l[4] for l in lines) or 1
print("total samples %d, total warp instructions %d" % (tot_s, tot_i))
print("--- by samples")
for f, n, src, s, i, st in sorted(lines, key=lambda l: -l[3])[:top]:
    ss = ",".join("%s=%d" % (k[6:], v) for k, v in sorted(st.items(), key=lambda kv: -kv[1])[:3])
    print("%5.1f%% smp %5.1f%% ins  %s:%d  %s   [%s]" % (100.0 * s / tot_s, 100.0 * i / tot_i, f, n, src[:90], ss))
print("--- by instructions")
for f, n, src, s, i, st in sorted(lines, key=lambda l: -l[4])[:top]:
    print("%5.1f%% ins %5.1f%% smp  %s:%d  %s" % (100.0 * i / tot_i, 100.0 * s / tot_s, f, n, src[:90]))
