"""SASS evidence for the round: per-kernel opcode histogram of aligator-swk_b200/libgarb200.so (cuobjdump -sass),
with the registers / spills ptxas reported (csrc/*.ptxas.log).  python tools/sass_summary.py r02 -> profiles/r02_sass_summary.txt
DMMA = mma.sync.m8n8k4.f64 (FP64 tensor pipe), UBLKCP = cp.async.bulk (1-D TMA), UBLKPF = bulk L2 prefetch,
SYNCS = mbarrier, LDL/STL = local-memory (spill) traffic, BAR = CTA / named barriers."""
import collections, glob, os, re, subprocess, sys

RR = sys.argv[1] if len(sys.argv) > 1 else "r02"
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
SO = os.path.join(ROOT, "aligator-swk_b200", "libgarb200.so")
out = subprocess.run(["cuobjdump", "-sass", SO], capture_output=True, text=True).stdout
kern, cur = collections.OrderedDict(), None
for line in out.splitlines():
    m = re.match(r"\s*Function : (\S+)", line)
    if m:
        cur = m.group(1); kern[cur] = collections.Counter(); continue
    m = re.match(r"\s+/\*[0-9a-f]+\*/\s+(?:@!?U?P\w+\s+)?([A-Z0-9_]+)", line)
    if m and cur:
        kern[cur][m.group(1)] += 1
regs = {}
for log in glob.glob(os.path.join(ROOT, "aligator-swk_b200", "csrc", "*.ptxas.log")):
    txt = open(log).read()
    for m in re.finditer(r"Compiling entry function '(\S+)' for 'sm_100a'.*?(\d+) bytes stack frame, (\d+) bytes spill stores, (\d+) bytes spill loads.*?Used (\d+) registers", txt, re.S):
        regs[m.group(1)] = (int(m.group(5)), int(m.group(3)), int(m.group(4)))
demangle = subprocess.run(["c++filt"], input="\n".join(kern), capture_output=True, text=True).stdout.splitlines()
cols = ["DMMA", "DFMA", "UBLKCP", "UBLKPF", "SYNCS", "LDS", "STS", "LDG", "STG", "LDL", "STL", "BAR", "SHFL"]
lines = [__doc__.strip(), "", "%-86s %6s %5s %11s " % ("kernel", "instr", "regs", "spill st/ld") + " ".join("%6s" % c for c in cols)]
tot = collections.Counter()
for (mangled, cnt), name in zip(kern.items(), demangle):
    name = re.sub(r"\(garb200::DevParams.*", "", name).replace("garb200::", "").replace("void ", "").replace("(int)", "").replace("(bool)", "")
    r = regs.get(mangled, (0, 0, 0))
    lines.append("%-86s %6d %5d %5d/%-5d " % (name[:86], sum(cnt.values()), r[0], r[1], r[2]) + " ".join("%6d" % cnt.get(c, 0) for c in cols))
    tot.update(cnt)
lines.append("%-86s %6d %5s %11s " % ("TOTAL (%d kernels)" % len(kern), sum(tot.values()), "", "") + " ".join("%6d" % tot.get(c, 0) for c in cols))
lines.append("")
lines.append("no tcgen05 / UTCMMA / HMMA / wgmma opcode anywhere: " + str(not any(k.startswith(("UTC", "HMMA", "HGMMA", "QGMMA")) for k in tot)))
open(os.path.join(ROOT, "profiles", f"{RR}_sass_summary.txt"), "w").write("\n".join(lines) + "\n")
print("\n".join(lines[-6:]))
