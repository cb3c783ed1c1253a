#!/usr/bin/env python
"""Instruction mix of a profiled kernel from an .ncu-rep (SASS page): executed warp-instructions by opcode.
Usage: python profiles/sass_mix.py gpurun_out/prof_bounce_c3.ncu-rep [kernel-index]"""
import collections, csv, io, subprocess, sys
rep = sys.argv[1]
which = int(sys.argv[2]) if len(sys.argv) > 2 else 0
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout
blocks = out.split('"Kernel Name",')[1:]
blk = blocks[which]
name, rest = blk.split("\n", 1)
rows = list(csv.reader(io.StringIO(rest)))
hdr = rows[0]
ii, it, isrc, ist = hdr.index("Instructions Executed"), hdr.index("Thread Instructions Executed"), hdr.index("Source"), hdr.index("# Samples")
agg = collections.Counter(); thr = collections.Counter(); smp = collections.Counter()
for r in rows[1:]:
    if len(r) <= it: continue
    toks = r[isrc].split()
    if not toks: continue
    op = toks[1] if toks[0].startswith("@") else toks[0]
    op = op.split(".")[0] + ("." + op.split(".")[1] if op.startswith(("LD", "ST", "MUFU")) and "." in op else "")
    agg[op] += int(r[ii]); thr[op] += int(r[it]); smp[op] += int(r[ist])
tot = sum(agg.values()); ts = sum(smp.values())
print(name[:120]); print(f"total warp-instructions {tot}, avg active threads {sum(thr.values())/tot:.1f}")
for op, n in agg.most_common(28):
    print(f"{op:14s} {n:12d} {100*n/tot:5.1f}%  thr/inst {thr[op]/max(n,1):5.1f}  stall-samples {100*smp[op]/ts:5.1f}%")
