#!/usr/bin/env python
"""Backward float-dataflow slices of a kernel's SASS (cuobjdump -sass output).

Used to learn the exact evaluation order (FFMA contraction, operand association) that nvcc/ptxas chose
for a kernel, so that bit-exact float parity can be engineered deliberately instead of hoped for.

    cuobjdump -sass foo.o > foo.sass
    python tools/sassflow.py foo.sass --kernel preprocess --exclude 0x1af0-0x4000 --stores
    python tools/sassflow.py foo.sass --kernel preprocess --at 0x4a10 --reg R12

For every selected value the tool walks backwards through the (linear) listing to the last writer of each
source register, skipping the excluded address ranges (untaken branch arms, slow-path subroutines), and
prints one line per instruction in dependency order:   t<addr> = OP(src, src, ...).
Non-float producers (loads, integer ops) are printed once as leaves with their raw operand text.
"""
from __future__ import annotations

import argparse
import re
import sys

INSTR = re.compile(r"^\s*/\*([0-9a-f]{4,5})\*/\s+(.*?);\s*/\*")
FLOAT_OPS = ("FFMA", "FMUL", "FADD", "MUFU", "FMNMX", "FSEL", "FSETP", "FCHK", "F2F", "I2F", "F2I", "FRND", "I2FP",
             "F2FP", "FMNMX3")
MOVES = ("MOV", "IMAD.MOV.U32", "IMAD.MOV")


class Ins:
    def __init__(self, addr, text):
        self.addr = addr
        self.text = text
        t = text.strip()
        self.pred = None
        if t.startswith("@"):
            p, t = t.split(None, 1)
            self.pred = p[1:]
        parts = t.split(None, 1)
        self.op = parts[0]
        ops = parts[1] if len(parts) > 1 else ""
        self.operands = [o.strip() for o in split_ops(ops)]
        self.base = self.op.split(".")[0]

    def dests(self):
        """registers written (names without modifiers)."""
        b = self.base
        if b in ("STG", "STS", "STL", "BRA", "EXIT", "BSSY", "BSYNC", "BAR", "RET", "CALL", "NOP", "RED", "ATOMG",
                 "WARPSYNC", "DEPBAR", "MEMBAR", "ERRBAR", "BREAK", "BPT", "YIELD", "ST"):
            return []
        if not self.operands:
            return []
        d = self.operands[0]
        out = []
        if b in ("FSETP", "ISETP", "PLOP3", "DSETP", "HSETP2", "FCHK", "R2P", "LOP3") and d.startswith(("P", "UP")):
            out.append(d)
            if len(self.operands) > 1 and re.fullmatch(r"U?P\d", self.operands[1]):
                out.append(self.operands[1])
            if b != "LOP3":
                return out
            # LOP3 with predicate dest also writes a register as operand 1
            d = self.operands[1]
        m = re.fullmatch(r"(U?R)(\d+)", d)
        if not m:
            return out
        n = int(m.group(2))
        width = 1
        if ".128" in self.op:
            width = 4
        elif ".64" in self.op or ".WIDE" in self.op or self.base in ("CS2R",):
            width = 2
        if self.base in ("LDC", "LDCU", "ULDC") and ".64" in self.op:
            width = 2
        out += [f"{m.group(1)}{n + i}" for i in range(width)]
        return out

    def sources(self):
        srcs = self.operands[1:] if self.dests() else self.operands
        return srcs


def split_ops(s):
    out, depth, cur = [], 0, ""
    for ch in s:
        if ch == "[":
            depth += 1
        elif ch == "]":
            depth -= 1
        if ch == "," and depth == 0:
            out.append(cur)
            cur = ""
        else:
            cur += ch
    if cur.strip():
        out.append(cur)
    return out


def parse(path, kernel):
    ins, active = [], False
    for line in open(path):
        if "Function :" in line:
            active = kernel in line
            if active and ins:
                break
            continue
        if not active:
            continue
        m = INSTR.match(line)
        if m:
            ins.append(Ins(int(m.group(1), 16), m.group(2)))
    return ins


REG = re.compile(r"(-?)(\|?)(U?R\d+|U?P\d+)(\.reuse|\.64|\.H0_H0|\.H1_H1)*(\|?)(\.reuse)?")


class Slicer:
    def __init__(self, ins, exclude):
        self.ins = ins
        self.exclude = exclude
        self.index = {i.addr: k for k, i in enumerate(ins)}
        self.names = {}
        self.lines = []

    def skipped(self, addr):
        return any(a <= addr < b for a, b in self.exclude)

    def writer(self, reg, before_idx):
        for k in range(before_idx - 1, -1, -1):
            i = self.ins[k]
            if self.skipped(i.addr):
                continue
            if reg in i.dests():
                return k
        return None

    def value(self, reg, before_idx, depth=0):
        """name of the value held by `reg` just before instruction index `before_idx`."""
        if reg in ("RZ", "URZ"):
            return "0"
        if reg in ("PT", "UPT"):
            return "true"
        k = self.writer(reg, before_idx)
        if k is None:
            return f"{reg}@entry"
        return self.emit(k, reg)

    def emit(self, k, reg):
        i = self.ins[k]
        key = (k, reg if len(i.dests()) > 1 else None)
        if key in self.names:
            return self.names[key]
        name = f"t{i.addr:04x}" + (f".{i.dests().index(reg)}" if len(i.dests()) > 1 else "")
        self.names[key] = name  # break cycles (loops)
        is_float = i.base in FLOAT_OPS or i.op in MOVES or i.base in ("HFMA2", "SEL", "LOP3", "SHF", "IADD3", "IMAD",
                                                                       "LEA", "PRMT")
        if i.op in MOVES or (i.base == "IMAD" and ".MOV" in i.op):
            src = i.operands[-1]
            expr = self.operand(src, k)
            if i.pred:
                old = self.value(reg, k)
                expr = f"({self.pred_expr(i.pred, k)} ? {expr} : {old})"
            else:
                # pure move: alias
                self.names[key] = expr
                return expr
        elif is_float:
            args = [self.operand(o, k) for o in i.sources()]
            expr = f"{i.op}({', '.join(args)})"
            if i.pred:
                old = self.value(reg, k)
                expr = f"({self.pred_expr(i.pred, k)} ? {expr} : {old})"
        else:
            expr = f"LEAF[{i.text.strip()}]"
        self.lines.append(f"{name} = {expr}")
        return name

    def pred_expr(self, p, k):
        neg = p.startswith("!")
        reg = p.lstrip("!")
        v = self.value(reg, k)
        return ("!" if neg else "") + v

    def operand(self, o, k):
        o = o.strip()
        if o.startswith(("c[", "desc[", "[")) or o.startswith("0x") or re.fullmatch(r"-?[\d.]+(e[+-]?\d+)?", o) or \
                o in ("+INF", "-INF", "+QNAN", "-QNAN"):
            return o
        m = REG.fullmatch(o)
        if m:
            neg, a1, reg, _, a2, _r = m.groups()
            v = self.value(reg, k)
            if a1:
                v = f"|{v}|"
            return ("-" if neg else "") + v
        if o.startswith("!"):
            return "!" + self.operand(o[1:], k)
        return o


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("sass")
    ap.add_argument("--kernel", required=True)
    ap.add_argument("--exclude", default="")
    ap.add_argument("--stores", action="store_true", help="slice every STG value")
    ap.add_argument("--at", default=None, help="address of an instruction: slice --reg before it")
    ap.add_argument("--reg", default=None)
    ap.add_argument("--range", default=None, help="only stores inside a-b")
    args = ap.parse_args()
    ins = parse(args.sass, args.kernel)
    if not ins:
        sys.exit("kernel not found")
    excl = []
    for r in filter(None, args.exclude.split(",")):
        a, b = r.split("-")
        excl.append((int(a, 16), int(b, 16)))
    sl = Slicer(ins, excl)
    lo, hi = 0, 1 << 30
    if args.range:
        a, b = args.range.split("-")
        lo, hi = int(a, 16), int(b, 16)
    if args.stores:
        for k, i in enumerate(ins):
            if i.base == "STG" and not sl.skipped(i.addr) and lo <= i.addr < hi:
                src = i.operands[-1]
                m = re.fullmatch(r"(R)(\d+)", src)
                width = 4 if ".128" in i.op else 2 if ".64" in i.op else 1
                vals = []
                if m:
                    for w in range(width):
                        vals.append(sl.value(f"R{int(m.group(2)) + w}", k))
                else:
                    vals.append(src)
                sl.lines.append(f"STORE@{i.addr:04x} {i.operands[0]} <- {', '.join(vals)}")
    if args.at:
        k = sl.index[int(args.at, 16)]
        sl.lines.append(f"VALUE {args.reg} before {args.at}: {sl.value(args.reg, k)}")
    print("\n".join(sl.lines))


if __name__ == "__main__":
    main()
