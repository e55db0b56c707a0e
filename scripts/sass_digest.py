#!/usr/bin/env python
"""Per-kernel digest of the tensor-core / TMA / TMEM opcodes in libtb_b200.so (cuobjdump -sass), written to
profiles/ so that the claim "tcgen05 + TMEM + TMA, no mma.sync" can be checked without disassembling anything.

    python scripts/sass_digest.py [out.txt]
"""
import collections
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = os.path.join(ROOT, "taboo_brittleness_b200", "libtb_b200.so")
PAT = re.compile(r"\b(UTC[A-Z]*MMA[.\w]*|LDTM[.\w]*|STTM[.\w]*|UTMALDG[.\w]*|UTMASTG[.\w]*|UTMAPF[.\w]*|UTCBAR[.\w]*|"
                 r"UTCATOMSWS[.\w]*|HMMA[.\w]*|IMMA[.\w]*|LDGSTS[.\w]*|MUFU\.[A-Z0-9]+|SYNCS[.\w]*|ELECT[.\w]*|"
                 r"UCGABAR[.\w]*|ATOMS?[.\w]*|RED[.\w]*)")


def main():
    out_path = sys.argv[1] if len(sys.argv) > 1 else os.path.join(ROOT, "profiles", "r02_sass_digest.txt")
    sass = subprocess.run(["cuobjdump", "-sass", LIB], capture_output=True, text=True, check=True).stdout
    names = subprocess.run(["cu++filt"], input="\n".join(re.findall(r"Function : (\S+)", sass)), capture_output=True,
                           text=True).stdout.splitlines()
    per = collections.OrderedDict()
    cur = None
    k = 0
    for line in sass.splitlines():
        m = re.search(r"Function : (\S+)", line)
        if m:
            cur = names[k] if k < len(names) else m.group(1)
            k += 1
            per[cur] = collections.Counter()
            per[cur]["(instructions)"] = 0
            continue
        if cur is None or "/*" not in line:
            continue
        body = line.split("/*")[1] if line.strip().startswith("/*") else line
        if re.match(r"\s*/\*[0-9a-f]{4}\*/", line):
            per[cur]["(instructions)"] += 1
            for op in PAT.findall(body):
                per[cur][op] += 1
    with open(out_path, "w") as f:
        f.write("# cuobjdump -sass taboo_brittleness_b200/libtb_b200.so -- opcode counts per kernel (sm_100a)\n"
                "# UTCHMMA = tcgen05.mma kind::f16, LDTM = tcgen05.ld, UTMALDG = cp.async.bulk.tensor (TMA load),\n"
                "# UTCBAR = tcgen05.commit, HMMA/IMMA = mma.sync (must not appear), LDGSTS = cp.async\n")
        total = collections.Counter()
        for name, c in per.items():
            short = name.split(">(")[0] + (">" if ">(" in name else "")
            short = short if ">(" in name else re.sub(r"\((?!int|bool).*", "", name)
            f.write(f"\n{short}\n")
            for op, n in sorted(c.items()):
                f.write(f"    {n:6d}  {op}\n")
                total[op] += n
        f.write("\n== whole library ==\n")
        for op, n in sorted(total.items()):
            f.write(f"    {n:6d}  {op}\n")
    print(out_path)


if __name__ == "__main__":
    main()
