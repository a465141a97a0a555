"""Dynamic SASS opcode mix of the first kernel in an `ncu --page source --csv` dump.
usage: ncu -i X.ncu-rep --page source --csv | python profiles/sass_mix.py [N_TOP]"""
import csv, sys, collections
rows = list(csv.reader(sys.stdin))
top = int(sys.argv[1]) if len(sys.argv) > 1 else 25
hdr = None
mix = collections.Counter()
samples = collections.Counter()
total = 0
kern = 0
for r in rows:
    if r and r[0] == "Kernel Name":
        kern += 1
        if kern > 1:
            break
        continue
    if r and r[0] == "Address":
        hdr = r
        continue
    if hdr is None or len(r) < len(hdr) - 5:
        continue
    src = r[hdr.index("Source")].strip()
    op = src.split()[0] if not src.startswith("@") else src.split()[1]
    op = op.rstrip(";").split(".")[0]
    n = int(r[hdr.index("Instructions Executed")] or 0)
    s = int(r[hdr.index("# Samples")] or 0)
    mix[op] += n
    samples[op] += s
    total += n
print("warp instructions executed:", total)
ALU = {"IADD3", "LOP3", "SHF", "PRMT", "LEA", "VIADD", "VIMNMX", "VIADDMNMX", "ISETP", "IABS", "MOV", "SEL", "IMNMX",
       "POPC", "FLO", "BREV", "SGXT", "BMSK", "PLOP3", "VABSDIFF", "VABSDIFF4", "FSEL", "FMNMX", "CS2R"}
FMA = {"IMAD", "FFMA", "FMUL", "FADD", "HFMA2", "IDP", "IDP4A"}
a = sum(v for k, v in mix.items() if k in ALU)
f = sum(v for k, v in mix.items() if k in FMA)
print("alu-pipe %.1f%%  fma-pipe %.1f%%  other %.1f%%" % (100 * a / total, 100 * f / total, 100 * (total - a - f) / total))
for op, n in mix.most_common(top):
    print("%-12s %10d  %5.1f%%   samples %6d" % (op, n, 100.0 * n / total, samples[op]))
