"""Per-source-line hot spots of an .ncu-rep captured with --import-source on (needs -lineinfo).

    python tools/ncu_lines.py gpurun_out/x.ncu-rep [top_n]
"""
import csv
import io
import subprocess
import sys


def main(path, top=14):
    raw = subprocess.run(["ncu", "-i", path, "--page", "source", "--print-source", "cuda,sass", "--csv"],
                         check=True, capture_output=True, text=True).stdout
    fn, hdr, agg = None, None, {}
    for r in csv.reader(io.StringIO(raw)):
        if not r:
            continue
        if r[0] == "File Path":
            fpath = r[1]
            continue
        if r[0] == "Function Name":
            fn = r[1]
            agg.setdefault(fn, {})
            continue
        if r[0] == "Line No":
            hdr = r
            i_s, i_i = hdr.index("# Samples"), hdr.index("Instructions Executed")
            continue
        if hdr and r[0] not in ("", "-") and r[0].isdigit():
            key = (fpath.split("/")[-1], int(r[0]), r[1].strip()[:90])
            try:
                s, ins = int(r[i_s]), int(r[i_i])
            except ValueError:
                continue
            a = agg[fn].setdefault(key, [0, 0])
            a[0] += s
            a[1] += ins
    for fn, lines in agg.items():
        ts = sum(v[0] for v in lines.values()) or 1
        ti = sum(v[1] for v in lines.values()) or 1
        print("\n== %s   samples=%d warp-instr=%d" % (fn[:80], ts, ti))
        for key, v in sorted(lines.items(), key=lambda kv: -kv[1][0])[:top]:
            print("  %5.1f%% smp %5.1f%% ins  %s:%d  %s" % (100 * v[0] / ts, 100 * v[1] / ti, key[0], key[1], key[2]))


if __name__ == "__main__":
    main(sys.argv[1], int(sys.argv[2]) if len(sys.argv) > 2 else 14)
