"""Source lines of one kernel sorted by executed warp instructions (ncu --import-source on report).
    python tools/ncu_ins.py report.ncu-rep kernel_substring [top_n]"""
import csv, io, subprocess, sys
raw = subprocess.run(["ncu", "-i", sys.argv[1], "--page", "source", "--print-source", "cuda,sass", "--csv"],
                     capture_output=True, text=True).stdout
agg, hdr, fn, f = {}, None, "", ""
for r in csv.reader(io.StringIO(raw)):
    if not r:
        continue
    if r[0] == "File Path":
        f = r[1].split('/')[-1]; continue
    if r[0] == "Function Name":
        fn = r[1]; continue
    if r[0] == "Line No":
        hdr = r; i_s, i_i = hdr.index("# Samples"), hdr.index("Instructions Executed"); continue
    if hdr and r[0].isdigit() and sys.argv[2] in fn:
        try:
            s, i = int(r[i_s]), int(r[i_i])
        except ValueError:
            continue
        a = agg.setdefault((f, int(r[0]), r[1].strip()[:90]), [0, 0]); a[0] += s; a[1] += i
ti = sum(v[1] for v in agg.values()) or 1
ts = sum(v[0] for v in agg.values()) or 1
print("total warp-instr", ti, "samples", ts)
for k, v in sorted(agg.items(), key=lambda kv: -kv[1][1])[:int(sys.argv[3]) if len(sys.argv) > 3 else 30]:
    print("%5.1f%% ins %5.1f%% smp %s:%d %s" % (100 * v[1] / ti, 100 * v[0] / ts, k[0], k[1], k[2]))
