"""Minimal SPIR-V reader + vectorised float32 evaluator for the reference's checked-in compiled shader
RayTracing/Assets/Shaders/Compiled/Ray/Test.spirv (the module Renderer.cs:356-357 loads; entry CSMain, :366).

TEST INFRASTRUCTURE ONLY (tests/ and tests/golden/make_spirv_golden.py).  This is the pin of the oracle to a
reference-held artefact: the module's words are decoded, its constants and operation list are asserted, and the
operation list is EVALUATED here, one numpy float32 operation per SPIR-V instruction, over whole pixel grids.

SPIR-V leaves the precision of four things to the implementation; this evaluator fixes them the way the frozen
oracle spec does (oracle/pt_oracle.h), and says so:
  OpFMul/OpFAdd/OpFSub/OpFDiv/OpFNegate   IEEE binary32, round to nearest even (numpy float32 arithmetic)
  OpDot                                   ((x0*y0 + x1*y1) + x2*y2), one rounding per operation
  GLSL.std.450 Sqrt                       correctly rounded
  GLSL.std.450 Normalize                  v / sqrt(dot(v, v)), three correctly rounded divisions
  GLSL.std.450 Fma                        a true fused multiply-add (one rounding), computed exactly through binary64
Only the opcodes that occur in Test.spirv are implemented; anything else raises.
"""
from __future__ import annotations

import hashlib
import struct
from dataclasses import dataclass, field
from typing import Dict, List, Optional, Tuple

import numpy as np

MAGIC = 0x07230203
TEST_SPIRV_SHA256 = "051e3943762e15c4331937cf14d8bcb8bb281fc8ba744d7a116d3593419c9f0f"

OP_NAMES = {
    3: "Source", 5: "Name", 11: "ExtInstImport", 12: "ExtInst", 14: "MemoryModel", 15: "EntryPoint", 16: "ExecutionMode",
    17: "Capability", 19: "TypeVoid", 20: "TypeBool", 21: "TypeInt", 22: "TypeFloat", 23: "TypeVector", 25: "TypeImage",
    32: "TypePointer", 33: "TypeFunction", 43: "Constant", 44: "ConstantComposite", 54: "Function", 56: "FunctionEnd",
    59: "Variable", 61: "Load", 71: "Decorate", 79: "VectorShuffle", 80: "CompositeConstruct", 81: "CompositeExtract",
    99: "ImageWrite", 112: "ConvertUToF", 127: "FNegate", 129: "FAdd", 131: "FSub", 133: "FMul", 136: "FDiv",
    142: "VectorTimesScalar", 148: "Dot", 186: "FOrdGreaterThan", 245: "Phi", 247: "SelectionMerge", 248: "Label",
    249: "Branch", 250: "BranchConditional", 253: "Return",
}
GLSL450 = {31: "Sqrt", 50: "Fma", 69: "Normalize"}


@dataclass
class Inst:
    op: int
    words: Tuple[int, ...]          # operands (without the first opcode/length word)

    @property
    def name(self) -> str:
        return OP_NAMES.get(self.op, f"Op{self.op}")


@dataclass
class Module:
    version: int
    generator: int
    bound: int
    insts: List[Inst]
    names: Dict[int, str] = field(default_factory=dict)

    def body(self) -> List[Inst]:
        """Instructions between OpFunction and OpFunctionEnd."""
        i0 = next(i for i, x in enumerate(self.insts) if x.op == 54)
        i1 = next(i for i, x in enumerate(self.insts) if x.op == 56)
        return self.insts[i0 + 1:i1]

    def float_constants(self) -> Dict[int, int]:
        """result id -> bit pattern of every 32-bit float OpConstant."""
        ftypes = {x.words[0] for x in self.insts if x.op == 22 and x.words[1] == 32}
        return {x.words[1]: x.words[2] for x in self.insts if x.op == 43 and x.words[0] in ftypes}

    def op_list(self) -> List[str]:
        """The arithmetic skeleton of the entry point, e.g. 'FMul', 'ExtInst:Normalize'."""
        out = []
        for x in self.body():
            if x.op == 12:
                out.append("ExtInst:" + GLSL450.get(x.words[3], str(x.words[3])))
            else:
                out.append(x.name)
        return out


def _string(words) -> str:
    raw = b"".join(struct.pack("<I", w) for w in words)
    return raw.split(b"\0", 1)[0].decode("utf-8", "replace")


def parse_words(words) -> Module:
    words = [int(w) for w in words]
    if words[0] != MAGIC:
        raise ValueError("not a SPIR-V module (bad magic)")
    m = Module(version=words[1], generator=words[2] >> 16, bound=words[3], insts=[])
    i = 5
    while i < len(words):
        op, n = words[i] & 0xFFFF, words[i] >> 16
        if n == 0 or i + n > len(words):
            raise ValueError("truncated instruction stream")
        m.insts.append(Inst(op, tuple(words[i + 1:i + n])))
        i += n
    for x in m.insts:
        if x.op == 5:
            m.names[x.words[0]] = _string(x.words[1:])
    return m


def parse_bytes(blob: bytes) -> Module:
    if len(blob) % 4:
        raise ValueError("SPIR-V size is not a multiple of 4")
    return parse_words(struct.unpack(f"<{len(blob) // 4}I", blob))


def sha256(blob: bytes) -> str:
    return hashlib.sha256(blob).hexdigest()


# ---- exact float32 fused multiply-add through binary64 -------------------------------------------------
def fma32(a, b, c):
    """RN32(a*b + c) for float32 arrays.  a*b is exact in binary64 (48-bit product); the sum is formed with an
    error-free TwoSum and rounded TO ODD in binary64, which makes the final rounding to binary32 the correct single
    rounding of the exact value (53 >= 2*24 + 2)."""
    a64, b64, c64 = (np.asarray(v, np.float32).astype(np.float64) for v in (a, b, c))
    p = a64 * b64                                   # exact
    s = p + c64
    bb = s - p
    err = (p - (s - bb)) + (c64 - bb)               # TwoSum: s + err == p + c exactly
    si = s.view(np.int64).copy()
    inexact = (err != 0) & np.isfinite(s)
    even = (si & 1) == 0
    # the exact value lies strictly between s and its neighbour in the direction of err: take whichever is odd
    toward_larger_magnitude = (err > 0) == (s > 0)
    adj = np.where(toward_larger_magnitude, 1, -1)
    zero = s == 0
    si = np.where(inexact & even & ~zero, si + adj, si)
    return si.view(np.float64).astype(np.float32)


def _dot(x, y):
    s = x[0] * y[0]
    for k in range(1, len(x)):
        s = s + x[k] * y[k]
    return s


def evaluate(m: Module, width: int, height: int, const_override: Optional[Dict[int, int]] = None):
    """Run the entry point for every invocation id.xy in [0,width) x [0,height).
    -> dict(image (H,W,4) float32 = the texels OpImageWrite stores, cond (H,W) bool = the branch condition,
            values = {result id: value}).  const_override: {float constant id: new bit pattern} (used to generalise
    the folded 1/1080 to another divisor; the pin itself runs the module unmodified)."""
    f32 = np.float32
    ys, xs = np.meshgrid(np.arange(height, dtype=np.uint32), np.arange(width, dtype=np.uint32), indexing="ij")
    val: Dict[int, object] = {}
    types: Dict[int, tuple] = {}
    builtin: Dict[int, int] = {}
    for x in m.insts:
        w = x.words
        if x.op == 71 and w[1] == 11:               # Decorate BuiltIn
            builtin[w[0]] = w[2]
        elif x.op == 22:
            types[w[0]] = ("float", w[1])
        elif x.op == 21:
            types[w[0]] = ("int", w[1], w[2])
        elif x.op == 20:
            types[w[0]] = ("bool",)
        elif x.op == 23:
            types[w[0]] = ("vec", w[1], w[2])
        elif x.op == 43:
            bits = (const_override or {}).get(w[1], w[2])
            t = types[w[0]]
            val[w[1]] = np.array(bits, np.uint32).view(f32)[()] if t[0] == "float" else np.uint32(bits)
        elif x.op == 44:
            val[w[1]] = [val[c] for c in w[2:]]
        elif x.op == 59:
            if builtin.get(w[1]) == 28:             # GlobalInvocationId
                val[w[1]] = [xs, ys, np.zeros_like(xs)]
            else:
                val[w[1]] = ("image", w[1])
    cond = None
    mask: Dict[int, object] = {}
    cur = None
    image = None
    with np.errstate(all="ignore"):
        for x in m.body():
            w = x.words
            if x.op == 248:
                cur = w[0]
            elif x.op == 61:
                val[w[1]] = val[w[2]]
            elif x.op == 79:
                src = list(val[w[2]]) + list(val[w[3]])
                val[w[1]] = [src[i] for i in w[4:]]
            elif x.op == 112:
                v = val[w[2]]
                val[w[1]] = [c.astype(f32) for c in v] if isinstance(v, list) else v.astype(f32)
            elif x.op in (133, 129, 131, 136):
                a, b = val[w[2]], val[w[3]]
                fn = {133: np.multiply, 129: np.add, 131: np.subtract, 136: np.divide}[x.op]
                if isinstance(a, list):
                    val[w[1]] = [fn(p, q, dtype=f32) for p, q in zip(a, b)]
                else:
                    val[w[1]] = fn(a, b, dtype=f32)
            elif x.op == 142:
                v, s = val[w[2]], val[w[3]]
                val[w[1]] = [np.multiply(c, s, dtype=f32) for c in v]
            elif x.op == 127:
                val[w[1]] = np.negative(val[w[2]], dtype=f32)
            elif x.op == 81:
                val[w[1]] = val[w[2]][w[3]]
            elif x.op == 80:
                val[w[1]] = [np.broadcast_to(np.asarray(val[c], f32), xs.shape) for c in w[2:]]
            elif x.op == 148:
                val[w[1]] = _dot(val[w[2]], val[w[3]])
            elif x.op == 12:
                kind = GLSL450.get(w[3])
                if kind == "Normalize":
                    v = [np.asarray(c, f32) for c in val[w[4]]]
                    ln = np.sqrt(_dot(v, v), dtype=f32)
                    val[w[1]] = [np.divide(c, ln, dtype=f32) for c in v]
                elif kind == "Sqrt":
                    val[w[1]] = np.sqrt(val[w[4]], dtype=f32)
                elif kind == "Fma":
                    val[w[1]] = fma32(val[w[4]], val[w[5]], val[w[6]])
                else:
                    raise NotImplementedError(f"GLSL.std.450 instruction {w[3]}")
            elif x.op == 186:
                val[w[1]] = np.asarray(val[w[2]], f32) > np.asarray(val[w[3]], f32)
            elif x.op == 247:
                pass
            elif x.op == 250:
                cond = val[w[0]]
                mask[w[1]] = cond
                mask[w[2]] = ~cond
            elif x.op == 249:
                pass
            elif x.op == 245:
                pairs = [(w[i], w[i + 1]) for i in range(2, len(w), 2)]
                (va, la), (vb, _lb) = pairs
                a, b = val[va], val[vb]
                val[w[1]] = [np.where(mask[la], p, q).astype(f32) for p, q in zip(a, b)]
            elif x.op == 99:
                image = np.stack([np.broadcast_to(np.asarray(c, f32), xs.shape) for c in val[w[2]]], axis=-1).astype(f32)
            elif x.op == 253:
                break
            else:
                raise NotImplementedError(f"SPIR-V opcode {x.op} ({x.name})")
    return {"image": image, "cond": np.asarray(cond, bool), "values": val}
