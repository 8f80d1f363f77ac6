#!/usr/bin/env python
"""Regenerate profiles/sass_evidence.txt and profiles/resource_usage.txt from the built library (no GPU needed):
per kernel, the counts of the SASS mnemonics that prove the hardware path (B200_PROFILING.md: tcgen05.mma -> UTC*MMA,
tcgen05.ld -> LDTM, TMA -> UTMALDG, tcgen05.commit -> UTCBAR, mbarrier -> SYNCS.*, vector reductions -> REDG.*.F32x4,
system-scope flag traffic -> ST/LD/ATOM *.STRONG.SYS, MEMBAR.*.SYS) and registers / stack / shared memory per kernel."""
import collections, os, re, subprocess
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
SO = os.path.join(ROOT, "ewdml", "_C", "libewdml.so")
KEEP = re.compile(r"\b(LDGMC\S*|UTC[A-Z]*MMA\S*|LDTM\S*|STTM\S*|UTMALDG\S*|UTMASTG\S*|UBLKCP\S*|UTCBAR\S*|UTCATOMSWS\S*|SYNCS\.\S+|REDG\.\S+|"
                  r"(?:ST|STG|LD|LDG|ATOMG?|RED)\.\S*STRONG\.SYS\S*|MEMBAR\.\S+|ATOMG\.\S+|ATOMS\.\S+|VOTEU?\.\S+|HMMA\S*|MUFU\.EX2)")


def demangle(names):
    out = subprocess.run(["c++filt"], input="\n".join(names), capture_output=True, text=True).stdout.splitlines()
    return [re.sub(r"\(.*$", "", re.sub(r"\(anonymous namespace\)::|<unnamed>::|^void ", "", o)) for o in out]


def main():
    sass = subprocess.run(["cuobjdump", "-sass", SO], capture_output=True, text=True).stdout
    per, cur = collections.OrderedDict(), None
    for line in sass.splitlines():
        m = re.search(r"Function : (\S+)", line)
        if m:
            cur = m.group(1)
            per[cur] = collections.Counter()
            continue
        if cur:
            for tok in KEEP.findall(line):
                per[cur][tok.rstrip(";")] += 1
    names = demangle(list(per))
    with open(os.path.join(ROOT, "profiles", "sass_evidence.txt"), "w") as f:
        f.write("# SASS evidence (cuobjdump -sass ewdml/_C/libewdml.so, sm_100a) -- mnemonic counts per kernel; scripts/sass_evidence.py\n\n"
                "tcgen05.mma -> UTC*MMA, tcgen05.ld -> LDTM, TMA load -> UTMALDG, tcgen05.commit -> UTCBAR, mbarrier -> SYNCS.*,\n"
                "red.global.add.v4.f32 -> REDG.E.ADD.F32x4, system-scope flag ops -> *.STRONG.SYS / MEMBAR.*.SYS,\n"
                "multimem.ld_reduce (NVLS in-switch reduction) -> LDGMC.E.ADD.F32x4; multimem.st is a plain STG to a multicast address\n"
                "(the PTX forms are listed in profiles/ptx_multimem.txt)\n\n")
        for name, c in sorted(zip(names, per.values()), key=lambda t: t[0]):
            if c:
                f.write(f"{name:58s} " + ", ".join(f"{k} x{v}" for k, v in sorted(c.items())) + "\n")
    res = subprocess.run(["cuobjdump", "-res-usage", SO], capture_output=True, text=True).stdout
    rows, fn = [], None
    for line in res.splitlines():
        m = re.search(r"Function (\S+):", line)
        if m:
            fn = m.group(1)
        m = re.search(r"REG:(\d+) STACK:(\d+) SHARED:(\d+) LOCAL:(\d+)", line)
        if m and fn:
            rows.append((fn, *map(int, m.groups())))
            fn = None
    names = demangle([r[0] for r in rows])
    with open(os.path.join(ROOT, "profiles", "resource_usage.txt"), "w") as f:
        f.write("# cuobjdump -res-usage ewdml/_C/libewdml.so: registers / stack bytes / static shared bytes per kernel (dynamic shared\n"
                "# memory of the GEMMs is set at launch: 3-8 stages x 24-48 KB); scripts/sass_evidence.py\n\n")
        for name, (_, reg, stack, shared, local) in sorted(zip(names, rows), key=lambda t: t[0]):
            f.write(f"{name:58s} REG {reg:3d}  STACK {stack:4d}  SHARED {shared:6d}  LOCAL {local}\n")
    print("kernels:", len(per), "with evidence:", sum(1 for c in per.values() if c))


if __name__ == "__main__":
    main()
