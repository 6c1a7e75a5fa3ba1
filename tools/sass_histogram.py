"""Per-kernel counts of the SASS mnemonics that prove which hardware paths the built library uses
(cuobjdump -sass ypp/libypp.so): UTCHMMA (tcgen05.mma), UTMALDG (TMA tensor load), LDTM (tcgen05.ld), UBLKCP (bulk copy),
SYNCS (mbarrier), plus HMMA/IMMA (legacy tensor-core paths: expected 0).  usage: python tools/sass_histogram.py > profiles/sass_histogram.txt"""
import collections, os, re, subprocess, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
lib = os.path.join(ROOT, "pytorch-object-detection_b200", "ypp", "libypp.so")
out = subprocess.run(["cuobjdump", "-sass", lib], capture_output=True, text=True, check=True).stdout
pats = ["UTCHMMA", "UTMALDG", "UTMASTG", "LDTM", "STTM", "UBLKCP", "SYNCS", "UTCBAR", "HMMA", "IMMA", "ACQBULK", "MATCH", "VOTE", "REDUX", "ATOMS", "BAR.SYNC"]
cur, table, total = None, collections.OrderedDict(), collections.Counter()
for line in out.splitlines():
    m = re.search(r"Function : (\S+)", line)
    if m:
        cur = m.group(1)
        table[cur] = collections.Counter()
        continue
    if cur is None:
        continue
    m = re.search(r"^\s+/\*[0-9a-f]+\*/\s+(?:@!?U?P\d+\s+)?([A-Z0-9_.]+)", line)
    if m:
        op = m.group(1)
        table[cur]["_n"] += 1
        for p in pats:
            if op.startswith(p):
                table[cur][p] += 1
arch = re.findall(r"arch = (sm_\w+)", out)
print("libypp.so: %d kernels, arch %s" % (len(table), sorted(set(arch))))
print("%-70s %7s " % ("kernel (demangled prefix)", "instrs") + " ".join("%8s" % p for p in pats))
for k, c in table.items():
    try:
        dem = subprocess.run(["c++filt", k], capture_output=True, text=True).stdout.strip()
    except Exception:
        dem = k
    m = re.search(r"((?:ypp|map|rs)_\w+(?:<[^>]*>)?)", dem)
    name = (m.group(1) if m else dem)[:70]
    print("%-70s %7d " % (name, c["_n"]) + " ".join("%8d" % c[p] for p in pats))
