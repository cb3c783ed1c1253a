#!/usr/bin/env python
"""Static opcode histogram of one kernel of the built library (cuobjdump -sass), whole kernel and per loop body.
Usage: python profiles/sass_static.py [mangled-name substring] > profiles/sass_csmain_accumulate.txt
A loop body = the instructions between a backward branch's target and the branch."""
import collections, os, re, subprocess, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
lib = os.path.join(ROOT, "pathtracing_b200", "libptb200.so")
want = sys.argv[1] if len(sys.argv) > 1 else "csmain_accumulateILi2ELb1ELb1"
out = subprocess.run(["cuobjdump", "-sass", lib], capture_output=True, text=True).stdout
fn, take, ins = None, False, []
for line in out.split("\n"):
    m = re.search(r"Function : (\S+)", line)
    if m:
        take = want in m.group(1)
        if take:
            fn = m.group(1)
        continue
    if take:
        m = re.match(r"\s+/\*([0-9a-f]{4})\*/\s+(.*?);", line)
        if m:
            ins.append((int(m.group(1), 16), m.group(2).strip()))
def opcode(t):
    toks = t.split()
    op = toks[1] if toks[0].startswith("@") else toks[0]
    return op.split(".")[0]
print(f"# cuobjdump -sass pathtracing_b200/libptb200.so, kernel {fn}")
print(f"# {len(ins)} SASS instructions.  FP32x2 = FFMA2 / FMUL2 / FADD2; a multiply that feeds an add is written as an FMA with a")
print("# run-time -0.0 addend (fmul2a), so no packed mul + add pair exists for ptxas to contract; scalar FMUL / FADD keep their .rn form.")
def hist(rows, title):
    c = collections.Counter(opcode(t) for _, t in rows)
    print(f"\n{title}: {len(rows)} instructions")
    print("  " + "  ".join(f"{k} {v}" for k, v in c.most_common()))
hist(ins, "whole kernel")
addr = {a: i for i, (a, _) in enumerate(ins)}
loops = []
for i, (a, t) in enumerate(ins):
    m = re.search(r"BRA(?:\.U)?\s+(?:!?U?P\d,\s*)?(?:!?U?P\d,\s*)?0x([0-9a-f]+)", t)
    if m and int(m.group(1), 16) < a and int(m.group(1), 16) in addr:
        loops.append((addr[int(m.group(1), 16)], i))
for n, (lo, hi) in enumerate(loops):
    body = ins[lo:hi + 1]
    c = collections.Counter(opcode(t) for _, t in body)
    kind = "FP32x2" if c["FFMA2"] + c["FMUL2"] + c["FADD2"] else "scalar"
    hist(body, f"loop {n} (0x{ins[lo][0]:04x}..0x{ins[hi][0]:04x}, {kind}: FFMA2 {c['FFMA2']} FMUL2 {c['FMUL2']} FADD2 {c['FADD2']})")
