#!/usr/bin/env python3
"""SASS opcode table of the product library (profiles/r2_sass_ops.md): python tools/sass_ops.py > profiles/r2_sass_ops.md
Static instruction counts per kernel from `cuobjdump -sass coloc_b200/libcoloc_b200.so`; needs no GPU."""
import os, re, subprocess, sys, collections
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
so = os.path.join(ROOT, "coloc_b200", "libcoloc_b200.so")
txt = subprocess.run(["cuobjdump", "-sass", so], capture_output=True, text=True, check=True).stdout
names = subprocess.run(["c++filt"], input="\n".join(re.findall(r"Function : (\S+)", txt)), capture_output=True, text=True).stdout.splitlines()
COLS = [("UTCOMMA", r"UTC\w*MMA"), ("LDTM", r"LDTM"), ("STTM", r"STTM"), ("UTCBAR", r"UTCBAR"), ("UBLKCP", r"UBLKCP"), ("UTMALDG", r"UTMALDG"),
        ("FFMA2", r"FFMA2"), ("FADD2", r"FADD2"), ("FMUL2", r"FMUL2"), ("POPC", r"POPC"), ("TLD", r"TLD"), ("IDP", r"IDP"),
        ("ACQBULK", r"ACQBULK"), ("PREEXIT", r"PREEXIT")]
rows = {}
for (mangled, body), name in zip(re.findall(r"Function : (\S+)\n(.*?)(?=\n\s*Function : |\Z)", txt, re.S), names):
    ops = re.findall(r"/\*[0-9a-f]{4}\*/\s+(?:@!?U?P\d+\s+)?([A-Z][A-Z0-9_.]*)", body)
    short = name.replace("(anonymous namespace)::", "").replace("(bool)", "").replace("(int)", "").replace("void ", "")
    short = re.sub(r"\(.*", "", short)
    c = collections.Counter()
    for col, pat in COLS:
        c[col] = sum(1 for o in ops if re.match(pat + r"(\.|$)", o))
    rows[short] = (len(ops), c)
print("# SASS opcode counts per kernel of the PRODUCT library (round 2)\n")
print("`python tools/sass_ops.py` = `cuobjdump -sass coloc_b200/libcoloc_b200.so` (sm_100a), static instruction counts.  UTC*MMA = tcgen05.mma, LDTM / STTM = "
      "tcgen05.ld / st, UTCBAR = tcgen05.commit, UBLKCP = cp.async.bulk (1-D TMA), UTMALDG = cp.async.bulk.tensor (tiled TMA), FFMA2 / FADD2 / FMUL2 = packed fp32, "
      "ACQBULK / PREEXIT = griddepcontrol.wait / launch_dependents (programmatic dependent launch).  The lab library (`make lab`, never loaded unless CLB_LIB=lab) "
      "additionally carries the superseded generations.\n")
print("| kernel | SASS instr | " + " | ".join(c for c, _ in COLS) + " |")
print("|---|---|" + "---|" * len(COLS))
for k in sorted(rows):
    n, c = rows[k]
    print(f"| `{k}` | {n} | " + " | ".join(str(c[col]) if c[col] else "" for col, _ in COLS) + " |")
n4 = rows.get("clb::match_tc4_kernel", (0,))[0]
print(f"\n`match_tc4_kernel`: {n4} instructions (round 1: 3 676 with the run-time `dbg` modes compiled in); no `printf`, no timing mode, no trace code in the product library.")
