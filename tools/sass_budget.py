#!/usr/bin/env python
"""Per-phase SASS instruction budget of a kernel in libmcsim_cuda.so.

    python tools/sass_budget.py <mangled-kernel-substring> [--csv out.csv] [--hist hist.csv]

The library is compiled with -lineinfo; `nvdisasm -gi` annotates every instruction with its inline stack.  Source
lines carry phase markers (`// @phase <name>` -- valid until the next marker or the end of the function); an
instruction belongs to the phase of the innermost frame of its stack that has one.  Counts are STATIC (one per
instruction in the binary): straight-line code of the unrolled step body executes once per step; loop bodies
(the SAT pair loops) are listed per iteration and have to be multiplied by the trip counts the kernel reports
(pair tests per step).  Pipes: FMA (FFMA/FMUL/FADD and the packed ...2 forms, IMAD), ALU (integer / logic / compare /
select / conversions that run there), XU (MUFU, I2F/F2I), LSU, UNIFORM (uniform datapath), CTRL (branches, barriers).
"""
import argparse
import collections
import os
import re
import subprocess
import sys
import tempfile

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
CSRC = os.path.join(ROOT, "probability-of-collision-for-safe-planning_b200", "csrc")
LIB = os.path.join(CSRC, "libmcsim_cuda.so")

PIPE = {}
for op in "FFMA FMUL FADD FFMA2 FMUL2 FADD2 IMAD HFMA2 IMAD.WIDE".split():
    PIPE[op] = "FMA"
for op in ("LOP3 IADD3 IADD ISETP FSETP SEL FSEL PRMT SHF LEA MOV FMNMX FMNMX3 IMNMX VIMNMX VIMNMX3 I2FP F2FP ISCADD POPC FLO BREV "
           "PLOP3 P2R R2P SGXT BMSK VOTE VOTEU CS2R S2R FCHK IABS LEA.HI FSET").split():
    PIPE[op] = "ALU"
for op in "MUFU I2F F2I F2F FRND".split():
    PIPE[op] = "XU"
for op in "LDS STS LDG STG LDL STL LDC LDCU ATOMS ATOMG REDG RED ATOM LDSM STSM REDUX SHFL MATCH".split():
    PIPE[op] = "LSU"
for op in "BRA BSSY BSYNC EXIT BAR WARPSYNC NOP CALL RET BREAK DEPBAR YIELD NANOSLEEP ERRBAR MEMBAR".split():
    PIPE[op] = "CTRL"


def pipe_of(op):
    base = op.split(".")[0]
    if base.startswith("U") and base not in ("UMOV_",) and base[1:] in ("MOV", "IADD3", "LEA", "ISETP", "IMAD", "LOP3", "SHF", "SEL", "POPC", "FLO", "PRMT", "FSETP", "FADD", "FMUL", "FFMA", "BREV", "PLOP3", "P2UR", "R2UR", "S2UR", "VOTEU", "IMNMX", "VIMNMX", "USEL", "F2FP", "I2FP"):
        return "UNIFORM"
    if base in ("S2UR", "R2UR", "VOTEU", "REDUX"):
        return "UNIFORM" if base != "REDUX" else "ALU"
    return PIPE.get(base, "OTHER")


def load_markers():
    """{file: [phase or None per line (1-based index)]}"""
    out = {}
    for fn in os.listdir(CSRC):
        if not fn.endswith((".cuh", ".h", ".cu")):
            continue
        lines = open(os.path.join(CSRC, fn)).read().split("\n")
        cur, depth, marks = None, 0, [None]
        for ln in lines:
            m = re.search(r"//\s*@phase\s+(\S+)", ln)
            if m:
                cur = m.group(1) if m.group(1) != "none" else None
            marks.append(cur)
            # a marker ends with the function it sits in: column-0 closing brace
            if ln.startswith("}"):
                cur = None
        out[fn] = marks
    return out


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("kernel")
    ap.add_argument("--lib", default=LIB)
    ap.add_argument("--csv")
    ap.add_argument("--hist")
    a = ap.parse_args()
    tmp = tempfile.mkdtemp()
    subprocess.run(["cuobjdump", "-xelf", "all", a.lib], cwd=tmp, check=True, capture_output=True)
    cub = [f for f in os.listdir(tmp) if f.endswith(".cubin") and "host_thread" not in f][0]
    dis = subprocess.run(["nvdisasm", "-gi", os.path.join(tmp, cub)], capture_output=True, text=True, check=True).stdout.split("\n")
    marks = load_markers()
    start = None
    for i, ln in enumerate(dis):
        if ln.startswith(".text.") and a.kernel in ln:
            start = i
            name = ln.strip().rstrip(":")[6:]
            break
    if start is None:
        sys.exit("kernel not found: " + a.kernel)
    phase_ops = collections.defaultdict(collections.Counter)
    hist = collections.Counter()
    stack = []
    fresh = False
    re_file = re.compile(r'//## File "([^"]+)", line (\d+)')
    re_ins = re.compile(r"^\s+/\*[0-9a-f]+\*/\s+(?:@!?U?P\w+\s+)?([A-Z][A-Z0-9_.]*)")
    for ln in dis[start + 1:]:
        if ln.startswith("//---") or (ln.startswith(".text.") and a.kernel not in ln):
            break
        m = re_file.search(ln)
        if m:
            if not fresh:
                stack, fresh = [], True
            stack.append((os.path.basename(m.group(1)), int(m.group(2))))
            continue
        m = re_ins.match(ln)
        if not m:
            continue
        fresh = False
        op = m.group(1)
        phase = "unmarked"
        for f, l in stack:          # innermost frame first
            mk = marks.get(f)
            if mk and l < len(mk) and mk[l]:
                phase = mk[l]
                break
        base = op.split(".")[0]
        if base == "IMAD" and ".WIDE" in op:
            base = "IMAD.WIDE"
        if base == "IMAD" and (".MOV" in op or ".SHL" in op or ".IADD" in op):
            base = "IMAD.MOV"   # integer moves / adds issued on the FMA pipe
            PIPE.setdefault(base, "FMA")
        phase_ops[phase][base] += 1
        hist[base] += 1
    pipes = ["FMA", "ALU", "XU", "LSU", "UNIFORM", "CTRL", "OTHER"]
    rows = []
    for ph in sorted(phase_ops, key=lambda p: -sum(phase_ops[p].values())):
        c = phase_ops[ph]
        by = collections.Counter()
        for op, n in c.items():
            by[pipe_of(op)] += n
        rows.append((ph, sum(c.values()), [by[p] for p in pipes], " ".join("%s:%d" % kv for kv in c.most_common(8))))
    print("kernel %s" % name)
    print("%-16s %6s  %s  top opcodes" % ("phase", "instr", " ".join("%7s" % p for p in pipes)))
    for ph, n, by, top in rows:
        print("%-16s %6d  %s  %s" % (ph, n, " ".join("%7d" % v for v in by), top))
    print("%-16s %6d" % ("total", sum(r[1] for r in rows)))
    if a.csv:
        with open(a.csv, "w") as f:
            f.write("kernel,%s\nphase,instructions,%s,top_opcodes\n" % (name, ",".join(pipes)))
            for ph, n, by, top in rows:
                f.write("%s,%d,%s,%s\n" % (ph, n, ",".join(str(v) for v in by), top))
    if a.hist:
        with open(a.hist, "w") as f:
            f.write("kernel,%s\nopcode,count,pipe\n" % name)
            for op, n in hist.most_common():
                f.write("%s,%d,%s\n" % (op, n, pipe_of(op)))


if __name__ == "__main__":
    main()
