"""Instruction budget of the trigger kernel's row body, from the SASS of the built object.

  python tools/k1_sass_table.py [--obj autobub3hs_b200/csrc/k1_trigger_tma.o] [--out profiles/r2_k1_sass.txt]

Finds k1_trigger_tma_kernel in `cuobjdump -sass`, locates the interior ring stage (the unchecked copy of do_stage: the
code between the mbarrier try-wait and the re-arming UTMALDG group whose rows carry no `rl` tests), splits it into its
T_R row bodies at the targets of the rows' skip branches, and prints for one interior row
  * the opcode histogram of the always-executed part (up to the branch over the rare 'a lane has a non-zero output' path)
  * the same for the rare path,
grouped by the phase of the algorithm the opcode can only belong to.  Also writes the full SASS of the kernel with the
row boundaries marked, so that the UTMALDG / VIADDMNMX / SYNCS lines can be seen in place."""
import argparse
import collections
import os
import re
import subprocess

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LINE = re.compile(r"^\s+/\*([0-9a-f]{4,})\*/\s+(?:@!?U?P\d+\s+)?([A-Z0-9_.]+)\s*(.*?);")


def kernel_sass(obj, name):
    txt = subprocess.run(["cuobjdump", "-sass", obj], capture_output=True, text=True, check=True).stdout
    out, on = [], False
    for l in txt.splitlines():
        if "Function :" in l:
            on = name in l
        if on:
            out.append(l)
    return out


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--obj", default=os.path.join(ROOT, "autobub3hs_b200", "csrc", "k1_trigger_tma.o"))
    ap.add_argument("--kernel", default="k1_trigger_tma_kernel")
    ap.add_argument("--out", default=os.path.join(ROOT, "profiles", "r2_k1_sass.txt"))
    a = ap.parse_args()
    lines = kernel_sass(a.obj, a.kernel)
    ins = []                                    # (addr, opcode, operands, predicated, raw)
    for l in lines:
        m = LINE.match(l)
        if m:
            ins.append((int(m.group(1), 16), m.group(2), m.group(3), "@" in l.split("*/")[1][:14], l))
    # the stage loop: first SYNCS.PHASECHK (try-wait) ... the back branch to it
    wait = next(i for i, x in enumerate(ins) if x[1].startswith("SYNCS.PHASECHK"))
    # row bodies of the unchecked stage: forward branches over the rare path land on the next row's first instruction
    tgt = collections.Counter()
    for addr, op, opr, _, _ in ins[wait:]:
        if op == "BRA":
            m = re.search(r"0x([0-9a-f]+)", opr)
            if m and int(m.group(1), 16) > addr:
                tgt[int(m.group(1), 16)] += 1
    starts = sorted(t for t, c in tgt.items() if c >= 2)            # each row end is the target of >= 2 skips
    # interior rows of the unchecked stage = the first run of equally long rows
    rows = [(starts[i], starts[i + 1]) for i in range(len(starts) - 1)]
    # the unchecked copy of the stage comes first in the code; its interior rows are the first two equally long rows
    s0, e0 = next((rows[i][0], rows[i][1]) for i in range(len(rows) - 1) if rows[i][1] - rows[i][0] == rows[i + 1][1] - rows[i + 1][0])
    body = [x for x in ins if s0 <= x[0] < e0]
    # always-executed part ends at the first branch whose target is the row end
    cut = next(i for i, x in enumerate(body) if x[1] == "BRA" and ("0x%x" % e0) in x[2].lower().replace("0x0", "0x"))
    hot, rare = body[:cut + 1], body[cut + 1:]

    def opname(op):
        p = op.split(".")
        if p[0] == "IMAD" and len(p) > 1 and p[1] in ("IADD", "MOV", "SHL", "U32", "WIDE", "HI"):
            return p[0] + "." + p[1]
        return p[0]

    def phase(op):
        if op.startswith("IMAD.IADD"): return "adds (IADD3 / IMAD.IADD / VIADD: the compiler spreads them over the ALU and FMA pipes)"
        if op.startswith("IMAD.MOV") or op.startswith("MOV") or op.startswith("UMOV"): return "moves"
        if op.startswith("LDS"): return "load 3 shared-memory vectors (frame j, frame j-d, K16)"
        if op.startswith("VIADDMNMX"): return "stage A: relu(+-(a-b) - t) on s16x2 (VIADDMNMX)"
        if op.startswith("PRMT"): return "PRMT: widen u8->s16x2 (8), odd-aligned pairs of the horizontal pass (10), pack bytes (4)"
        if op.startswith("IADD3") or op.startswith("IADD") or op.startswith("VIADD"): return "adds (IADD3 / IMAD.IADD / VIADD: the compiler spreads them over the ALU and FMA pipes)"
        if op.startswith("IMAD") or op.startswith("LEA") or op.startswith("SHF") or op.startswith("LOP3"): return "multiply-adds of the horizontal pass: e + 6 W + 4 t (IMAD / LEA)"
        if op.startswith("SHFL"): return "SHFL: halo words from the neighbour lanes (4)"
        if op.startswith("VABSDIFF"): return "|P - N| per byte (VABSDIFF4)"
        if op.startswith("VOTE") or op.startswith("ISETP") or op.startswith("BRA") or op.startswith("PLOP") or op.startswith("POPC") or op.startswith("SEL"): return "hit vote + branch over the rare path"
        return "other"
    os.makedirs(os.path.dirname(a.out), exist_ok=True)
    with open(a.out, "w") as f:
        f.write("# k1_trigger_tma_kernel (autobub3hs_b200/csrc/k1_trigger_tma.cu), sm_100a SASS, written by tools/k1_sass_table.py\n")
        f.write("# one interior row of the unchecked ring stage = addresses 0x%04x .. 0x%04x: %d instructions always executed, %d on the\n"
                "# rare path (a lane of the warp has a non-zero output: append to the pending list, every 32 entries the byte-wise\n"
                "# shared-memory histogram).  One row = 240 output pixels (30 lanes x 8) of one frame pair.\n#\n" % (s0, e0, len(hot), len(rare)))
        for title, part in (("always executed", hot), ("rare path", rare)):
            f.write("# %s: %d instructions\n" % (title, len(part)))
            byop = collections.Counter(opname(x[1]) for x in part)
            byph = collections.defaultdict(int)
            for x in part:
                byph[phase(x[1])] += 1
            for op, c in byop.most_common():
                f.write("#   %-12s %3d\n" % (op, c))
            f.write("#  by phase:\n")
            for ph, c in sorted(byph.items(), key=lambda t: -t[1]):
                f.write("#   %3d  %s\n" % (c, ph))
            f.write("#\n")
        tma = [x for x in ins if x[1].startswith("UTMALDG")]
        f.write("# TMA: %d UTMALDG sites (%s); mbarrier: %d SYNCS sites; VIADDMNMX sites: %d; register count and spills: see -Xptxas -v below\n#\n"
                % (len(tma), ", ".join(sorted({x[1] for x in tma})), sum(1 for x in ins if x[1].startswith("SYNCS")),
                   sum(1 for x in ins if x[1].startswith("VIADDMNMX"))))
        for l in lines:
            m = LINE.match(l)
            if m and int(m.group(1), 16) == s0:
                f.write("        // ======== interior row body starts (tools/k1_sass_table.py) ========\n")
            if m and int(m.group(1), 16) == e0:
                f.write("        // ======== interior row body ends ========\n")
            if re.match(r"^\s+/\* 0x[0-9a-f]+ \*/\s*$", l):        # second encoding word: drop
                continue
            f.write(re.sub(r"\s+/\* 0x[0-9a-f]+ \*/\s*$", "", l) + "\n")
    print(open(a.out).read().split("        //")[0][:4000])


if __name__ == "__main__":
    main()
