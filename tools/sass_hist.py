"""Opcode histogram of one steady-state step (between two BAR.SYNCs) of each k_sweep kernel in a binary."""
import re, sys, collections, subprocess
txt = subprocess.run(["cuobjdump", "-sass", sys.argv[1]], capture_output=True, text=True).stdout
funcs = re.split(r"\n\s+Function : ", txt)[1:]
ALU = ("VIADDMNMX", "VIMNMX", "VIMNMX3", "PRMT", "IADD3", "LOP3", "SEL", "SHF", "ISETP", "VIADD", "PLOP3", "LEA", "IABS", "FLO", "POPC", "BREV", "IMNMX", "MOV")
for f in funcs:
    name = f.split("\n")[0]
    if len(sys.argv) > 2 and sys.argv[2] not in name: continue
    ops = [m.group(2) for m in re.finditer(r"/\*[0-9a-f]{4,5}\*/\s+(@!?U?P\d\s+)?([A-Z0-9_.]+)", f)]
    bars = [i for i, o in enumerate(ops) if o.startswith("BAR")]
    segs = [(bars[i + 1] - bars[i], bars[i], bars[i + 1]) for i in range(len(bars) - 1)]
    big = sorted(s for s in segs if s[0] > 100)
    if not big: continue
    s = big[len(big) // 2]
    c = collections.Counter()
    for o in ops[s[1]:s[2]]:
        parts = o.split(".")
        key = parts[0] + ("." + parts[1] if len(parts) > 1 and parts[0] in ("IMAD", "LDS", "STS", "VIMNMX", "VIMNMX3", "VIADDMNMX") else "")
        c[key] += 1
    alu = sum(v for k, v in c.items() if k.split(".")[0] in ALU)
    fma = sum(v for k, v in c.items() if k.startswith("IMAD"))
    print(f"{name}\n  step {s[0]} instr: alu-pipe {alu}, fma-pipe {fma}; segs {[x[0] for x in segs]}")
    print("  " + " ".join(f"{k}:{v}" for k, v in sorted(c.items(), key=lambda x: -x[1])))
