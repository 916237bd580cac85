"""Developer tool: aggregate an ncu report's warp-stall samples by CUDA source line for one kernel.
usage: python tools/ncu_lines.py report.ncu-rep kernel-regex [min-percent]"""
import csv, subprocess, sys, io
rep, kern = sys.argv[1], sys.argv[2]
minpct = float(sys.argv[3]) if len(sys.argv) > 3 else 1.0
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "cuda,sass", "--kernel-name", "regex:" + kern],
                     capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(out)))
hdr = None; cur_file = ""; agg = {}; tot = 0
for r in rows:
    if len(r) >= 2 and r[0] == "File Path": cur_file = r[1].split("/")[-1]; continue
    if r and r[0] == "Line No": hdr = r; continue
    if hdr is None or not r or not r[0].strip().isdigit(): continue
    i_s = hdr.index("# Samples"); i_e = hdr.index("Instructions Executed")
    s = int(r[i_s]) if r[i_s].isdigit() else 0
    e = int(r[i_e]) if r[i_e].isdigit() else 0
    k = (cur_file, int(r[0]), r[1].strip()[:100])
    a = agg.setdefault(k, [0, 0]); a[0] += s; a[1] += e; tot += s
print("total samples", tot)
for (f, ln, src), (s, e) in sorted(agg.items()):
    if s >= tot * minpct / 100: print(f"{f}:{ln:4d} {100*s/tot:5.1f}% inst={e:9d}  {src}")
