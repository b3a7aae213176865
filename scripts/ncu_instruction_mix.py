#!/usr/bin/env python
"""Executed warp instructions and stall samples by instruction class, from the SASS source page of an .ncu-rep:
python scripts/ncu_instruction_mix.py report.ncu-rep "header comment" > profiles/xyz_instruction_mix.txt"""
import csv
import re
import subprocess
import sys

rep, note = sys.argv[1], sys.argv[2] if len(sys.argv) > 2 else ""
raw = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(raw.splitlines()))
h = next(r for r in rows if "Source" in r and "# Samples" in r)
data = rows[rows.index(h) + 1:]
i_src, i_smp, i_ex = h.index("Source"), h.index("# Samples"), h.index("Instructions Executed")

CLASSES = [
    ("FMNMX / FMNMX3 (min / max trees)", r"^FMNMX"),
    ("mbarrier try_wait (polls)", r"^SYNCS\.PHASECHK|^NANOSLEEP"),
    ("mbarrier arrive / expect_tx", r"^SYNCS\.(ARRIVE|EXCH|CCTL)"),
    ("LDTM (tcgen05.ld)", r"^LDTM"),
    ("tcgen05.mma / commit / TMA bulk / tcgen05.st", r"^(UTCHMMA|UTCBAR|UBLKCP|STTM|UTCATOMSWS|UCGABAR|FENCE)"),
    ("shared memory loads / stores", r"^(LDS|STS|LDSM|ATOMS)"),
    ("global / local / constant memory", r"^(LDG|STG|LDL|STL|LDC|LDCU|ATOMG|RED|LD\b|ST\b)"),
    ("compares / predicates", r"^(FSETP|ISETP|PLOP3|PSETP|UISETP|UPLOP3|SEL|FSEL|USEL|VOTE|VOTEU|P2R|R2P)"),
    ("branches / sync / nop", r"^(BRA|BRX|BSSY|BSYNC|WARPSYNC|NOP|YIELD|EXIT|CALL|RET|BAR|BREAK|ELECT|JMP|BPT|DEPBAR|ERRBAR|MEMBAR|BMOV)"),
]
tot_i = tot_s = 0
acc = {name: [0, 0] for name, _ in CLASSES}
acc["integer / move / other"] = [0, 0]
for r in data:
    if len(r) <= max(i_src, i_smp, i_ex):
        continue
    op = re.sub(r"^@!?U?P\d+\s+", "", r[i_src].strip())
    ex, sm = int(r[i_ex] or 0), int(r[i_smp] or 0)
    tot_i += ex
    tot_s += sm
    for name, pat in CLASSES:
        if re.match(pat, op):
            acc[name][0] += ex
            acc[name][1] += sm
            break
    else:
        acc["integer / move / other"][0] += ex
        acc["integer / move / other"][1] += sm
print(f"# {note}")
print("# executed warp instructions by class (ncu source page of the SAME capture as the summary csv)")
print(f"total executed warp instructions: {tot_i}")
print(f"total stall samples: {tot_s}\n")
print(f"{'class':56s} {'instructions':>13s} {'share':>7s} {'samples':>9s} {'share':>7s}")
for name, (ex, sm) in sorted(acc.items(), key=lambda kv: -kv[1][0]):
    print(f"{name:56s} {ex:13d} {100.0 * ex / max(tot_i, 1):6.1f}% {sm:9d} {100.0 * sm / max(tot_s, 1):6.1f}%")
