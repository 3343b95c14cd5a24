#!/usr/bin/env python3
"""Instruction mix of the hot loop of a kernel from `cuobjdump -sass` output (stdin or file).
The hot loop is taken as the backward branch whose body has the highest FP-instruction density."""
import re, sys, collections
txt = open(sys.argv[1]).read() if len(sys.argv) > 1 else sys.stdin.read()
ins = []
for line in txt.splitlines():
    m = re.match(r"\s+/\*([0-9a-f]{4,})\*/\s+(.*?);", line)
    if m:
        ins.append((int(m.group(1), 16), m.group(2).strip()))
addr_idx = {a: k for k, (a, _) in enumerate(ins)}
best = None
for k, (a, t) in enumerate(ins):
    m = re.search(r"BRA\S*\s+(?:\S+,\s*)?0x([0-9a-f]+)", t)
    if m:
        tgt = int(m.group(1), 16)
        if tgt < a and tgt in addr_idx:
            body = ins[addr_idx[tgt]:k + 1]
            nfp = sum(1 for _, x in body if re.search(r"\b(FMUL2?|FADD2?|FFMA2?|DMUL|DADD|DFMA)\b", x))
            dens = nfp / max(1, len(body))
            if nfp >= 8 and (best is None or dens > best[0]):
                best = (dens, body)
if not best:
    sys.exit("no loop found")
body = best[1]
cnt = collections.Counter()
for _, x in body:
    x = re.sub(r"^@!?U?P\d+\s+", "", x)
    cnt[x.split()[0].split(".")[0]] += 1
print("loop: %d instructions, 0x%x..0x%x" % (len(body), body[0][0], body[-1][0]))
for k, v in cnt.most_common():
    print("  %-10s %d" % (k, v))
