#!/usr/bin/env python
"""SASS evidence for libomplcu.so: python scripts/sass_excerpt.py > profiles/r2_sass_excerpt.txt
Per-kernel counts of the instructions that prove tcgen05 / TMEM / TMA / cluster multicast / redux use, and excerpts
of the headline kernel's MMA issue and selection step."""
import collections
import os
import re
import subprocess

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = os.path.join(ROOT, "ompl_b200", "lib", "libomplcu.so")
sass = subprocess.run(["cuobjdump", "-sass", LIB], capture_output=True, text=True).stdout
arch = sorted(set(re.findall(r"arch = (sm_\w+)", sass)))
COLS = [("UTCHMMA", r"\bUTCHMMA\b"), ("LDTM", r"\bLDTM"), ("STTM", r"\bSTTM"), ("UTCBAR", r"\bUTCBAR\b(?!\.MULTI)"),
        ("UTCBAR.MC", r"UTCBAR\.MULTICAST"), ("UBLKCP", r"UBLKCP\.S\.G(?!\.MULTI)"), ("UBLKCP.MC", r"UBLKCP\.S\.G\.MULTICAST"),
        ("UCGABAR", r"UCGABAR"), ("TRYWAIT", r"TRYWAIT"), ("REDUX", r"\bREDUX"), ("FFMA2", r"\bFFMA2\b"), ("FMNMX3", r"\bFMNMX3\b")]
kern = collections.OrderedDict()
cur = None
for line in sass.splitlines():
    m = re.match(r"\s*Function : (\S+)", line)
    if m:
        cur = m.group(1)
        kern[cur] = []
    elif cur is not None:
        kern[cur].append(line)
names = subprocess.run(["cu++filt"], input="\n".join(kern), capture_output=True, text=True).stdout.splitlines()
short = lambda s: re.sub(r"\(anonymous namespace\)::|omplcu::|<unnamed>::|\(int\)|\(bool\)", "", s)  # noqa: E731
print("# SASS evidence for libomplcu.so (cuobjdump -sass ompl_b200/lib/libomplcu.so; scripts/sass_excerpt.py)")
print("# tcgen05.mma -> UTCHMMA, tcgen05.ld / st -> LDTM / STTM, tcgen05.commit -> UTCBAR (.MULTICAST: to both CTAs of a pair),")
print("# cp.async.bulk -> UBLKCP (.MULTICAST: into both CTAs' rings), barrier.cluster -> UCGABAR, mbarrier try_wait -> TRYWAIT,")
print("# redux.sync -> REDUX, fma.rn.f32x2 -> FFMA2")
print(f"\narchitectures in the library: {', '.join(arch)}\n")
print(f"{'kernel':92s} " + " ".join(f"{c:>9s}" for c, _ in COLS))
tot = collections.Counter()
others = collections.Counter()
nother = 0
for (mangled, body), name in zip(kern.items(), names):
    txt = "\n".join(body)
    cnt = {c: len(re.findall(p, txt)) for c, p in COLS}
    tot.update(cnt)
    if any(cnt[c] for c in ("UTCHMMA", "LDTM", "UBLKCP.MC")) or cnt["REDUX"] >= 3:
        print(f"{short(name)[:92]:92s} " + " ".join(f"{cnt[c]:9d}" for c, _ in COLS))
    else:
        others.update(cnt)
        nother += 1
print(f"{'(' + str(nother) + ' other kernels: FP32 scans, re-rank, validity, ...)':92s} " + " ".join(f"{others[c]:9d}" for c, _ in COLS))
print(f"{'whole library':92s} " + " ".join(f"{tot[c]:9d}" for c, _ in COLS))
print(f"HMMA / wgmma instructions in the library: {len(re.findall(r'HMMA', sass)) - tot['UTCHMMA']}")


def excerpt(title, pred, pat, before, after):
    for (mangled, body), name in zip(kern.items(), names):
        if pred(name):
            for i, l in enumerate(body):
                if re.search(pat, l):
                    print(f"\n## {title}\n## {short(name)}")
                    for x in body[max(0, i - before):i + after]:
                        if "/*" in x and not re.match(r"\s*/\* 0x", x):
                            print(re.sub(r"\s*/\* 0x[0-9a-f]+ \*/\s*$", "", x))
                    return


d7 = lambda n: "knn_scan_tc_kernel<16, 4, 0, 0, 0>" in short(n)  # noqa: E731
excerpt("MMA issue of the headline d = 7 kernel (two tcgen05.mma of K = 16 per 64-column half tile, commit)", d7, r"UTCHMMA", 14, 22)
excerpt("selection step of the same kernel: tcgen05.ld of 32 columns, FMNMX3 min-tree, threshold test", d7, r"LDTM", 4, 26)
excerpt("CTA pairs (d >= 53): multicast bulk copy of half a point chunk into both rings",
        lambda n: "knn_scan_tc_kernel<16, 1, 0, 1, 1>" in short(n), r"UBLKCP\.S\.G\.MULTICAST", 8, 4)
excerpt("CTA pairs: tcgen05.commit multicast on the ring's empty barrier",
        lambda n: "knn_scan_tc_kernel<16, 1, 0, 1, 1>" in short(n), r"UTCBAR\.MULTICAST", 6, 3)
excerpt("few-query scan: redux.sync minimum over the batch / maximum over the list",
        lambda n: "knn_few_kernel<1>" in short(n), r"REDUX", 6, 14)
