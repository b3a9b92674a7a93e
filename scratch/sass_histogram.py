"""opcode histogram per kernel of libb2nn.so (cuobjdump -sass): what proves the kernels are Blackwell-native
(UTC*MMA = tcgen05.mma, LDTM = tcgen05.ld, UTMALDG = cp.async.bulk.tensor, UBLKCP = cp.async.bulk, LDGMC/STGMC... =
multimem).  Usage: python scratch/sass_histogram.py > profiles/r02_sass_histogram.txt"""
import collections, os, re, subprocess, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
lib = os.path.join(ROOT, "graph-neural-net_b200", "lib", "libb2nn.so")
out = subprocess.run(["cuobjdump", "-sass", lib], capture_output=True, text=True).stdout
res = subprocess.run(["cuobjdump", "--dump-resource-usage", lib], capture_output=True, text=True).stdout
usage = {}
for m in re.finditer(r"Function (\S+):\s*\n\s*(REG:\d+.*)", res):
    usage[m.group(1)] = m.group(2)
kern, hist = None, {}
for line in out.splitlines():
    m = re.match(r"\s*Function : (\S+)", line)
    if m:
        kern = m.group(1)
        hist[kern] = collections.Counter()
        continue
    m = re.match(r"\s*/\*[0-9a-f]{4,6}\*/\s+(?:@!?U?P\d+\s+)?([A-Z][A-Z0-9_.]*)", line)
    if m and kern:
        hist[kern][m.group(1)] += 1
demangle = subprocess.run(["c++filt"] + list(hist), capture_output=True, text=True).stdout.splitlines()
KEY = ("UTC", "LDTM", "STTM", "UTMA", "UBLKCP", "LDGMC", "STGMC", "REDGMC", "MULTIMEM", "SYNCS", "HMMA", "DMMA", "DFMA", "DADD", "DMUL",
       "LDGSTS", "ATOM", "RED", "CCTL", "ELECT", "UTCBAR", "ACQBULK", "MEMBAR", "ERRBAR")
print("# cuobjdump -sass of graph-neural-net_b200/lib/libb2nn.so (sm_100a): instructions per kernel, Blackwell-specific and")
print("# tensor / FP64 / atomic opcodes listed by full mnemonic, the rest summed by base opcode (top 12)")
for name, dm in sorted(zip(hist, demangle), key=lambda p: p[1]):
    h = hist[name]
    tot = sum(h.values())
    print("\n== %s\n   %d instructions; %s" % (dm[:200], tot, usage.get(name, "")))
    special = {k: v for k, v in h.items() if k.startswith(KEY)}
    if special:
        print("   " + ", ".join("%s x%d" % kv for kv in sorted(special.items())))
    base = collections.Counter()
    for k, v in h.items():
        base[k.split(".")[0]] += v
    print("   top: " + ", ".join("%s %d" % kv for kv in base.most_common(12)))
