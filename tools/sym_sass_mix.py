#!/usr/bin/env python
"""Instruction mix of the pair-symmetric kernels' hot loop (the unclamped strip loop: the first
backward-branch loop that contains SHFL), per pair evaluation, from `cuobjdump -sass` of the built
library.  Companion of tools/sass_mix.py (default kernel).  Usage: python tools/sym_sass_mix.py"""
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = os.path.join(ROOT, "parallel_comp_labs_b200", "libnbody_b200.so")


def functions():
    out = subprocess.run(["cuobjdump", "-sass", LIB], capture_output=True, text=True, check=True).stdout
    fn, d = None, {}
    for ln in out.splitlines():
        m = re.search(r"Function : (\S+)", ln)
        if m:
            fn = m.group(1)
            d[fn] = []
            continue
        m = re.match(r"\s+/\*([0-9a-f]{4,})\*/\s+(.*?)\s*;", ln)
        if m and fn:
            d[fn].append((int(m.group(1), 16), m.group(2)))
    return d


def hot_loop(ins):
    loops = []
    for k, (addr, text) in enumerate(ins):
        m = re.search(r"BRA(?:\.U)?\s+(?:!?U?P\d,\s+)?0x([0-9a-f]+)", text)
        if m and int(m.group(1), 16) < addr:
            start = int(m.group(1), 16)
            body = [t for a, t in ins if start <= a <= addr]
            if any("SHFL" in t for t in body) and any("DFMA" in t or "FFMA" in t for t in body):
                loops.append((start, addr, body))
    # innermost loops only (no other loop nested inside), the first of them in program order
    inner = [l for l in loops if not any(o is not l and l[0] <= o[0] and o[1] <= l[1] for o in loops)]
    return min(inner)[2] if inner else None


def main():
    rows = []
    for name, ins in sorted(functions().items()):
        m = re.search(r"force_sym_kernelI([df])Li(\d+)ELi(\d+)ELi(\d+)ELi(\d+)ELi(\d+)ELi(\d+)ELi(\d+)E", name)
        if not m:
            continue
        t, th, i, st, mb, um, order, rot = m.group(1), *map(int, m.groups()[1:])
        body = hot_loop(ins)
        if body is None:
            continue
        ops = {}
        for text in body:
            op = text.split()[1] if text.startswith("@") else text.split()[0]
            op = op.split(".")[0]
            ops[op] = ops.get(op, 0) + 1
        mufu = ops.get("MUFU", 0)                 # one per pair evaluation
        fp = sum(v for k, v in ops.items() if k in (("DFMA", "DMUL", "DADD") if t == "d" else
                                                     ("FFMA", "FMUL", "FADD", "FFMA2", "FMUL2", "FADD2")))
        other = sum(ops.values()) - fp
        rows.append((t, i, mb, um, rot, fp / mufu, ops.get("SHFL", 0) / mufu, other / mufu, len(body), mufu))
    print("type I minCTA UM ROT | FP-pipe instr/pair-eval | SHFL/pair-eval | all other/pair-eval | loop instrs (pair-evals per trip)")
    for r in rows:
        print("%s    %d  %d      %d  %d   | %6.2f | %5.2f | %5.2f | %d (%d)" % r)


if __name__ == "__main__":
    sys.exit(main())
