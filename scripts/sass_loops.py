"""Lists the backward-branch loops of one kernel in a cuobjdump -sass dump with their instruction mix.
usage: cuobjdump -sass x.o | python scripts/sass_loops.py <kernel-name-substring> [min LDS.U8 count]"""
import re, sys, collections
name = sys.argv[1]
min_lds = int(sys.argv[2]) if len(sys.argv) > 2 else 8
ins, on = [], False
for line in sys.stdin:
    if "Function :" in line:
        on = name in line
        continue
    if not on:
        continue
    m = re.match(r"\s+/\*([0-9a-f]{4,})\*/\s+(.*?);", line)
    if m:
        ins.append((int(m.group(1), 16), m.group(2).strip()))
addr = {a: i for i, (a, _) in enumerate(ins)}
print("kernel instructions:", len(ins))
for i, (a, t) in enumerate(ins):
    m = re.search(r"\bBRA(?:\.U)?\s+(?:\S+,\s*)?(0x[0-9a-f]+)", t)
    if not m:
        continue
    tgt = int(m.group(1), 16)
    if tgt >= a or tgt not in addr:
        continue
    body = ins[addr[tgt]: i + 1]
    mix = collections.Counter()
    for _, x in body:
        op = x.split()[1] if x.startswith("@") else x.split()[0]
        mix[op.split(".")[0] + ("." + op.split(".")[1] if op.startswith(("IMAD", "LDS", "SHF", "LEA")) and "." in op else "")] += 1
    if mix["LDS.U8"] < min_lds:
        continue
    print("loop 0x%x..0x%x: %d instructions, %d LDS.U8 -> %.2f per cell" % (tgt, a, len(body), mix["LDS.U8"], len(body) / mix["LDS.U8"]))
    print("   " + ", ".join("%s %d" % kv for kv in mix.most_common(14)))
