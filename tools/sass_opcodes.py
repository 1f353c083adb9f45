#!/usr/bin/env python3
"""Per-kernel counts of the SASS mnemonics that prove a Blackwell-native path (B200_PROFILING.md: tcgen05.mma -> UTC*MMA,
tcgen05.ld/st -> LDTM/STTM, tcgen05.cp -> UTCCP, TMA -> UTMALDG/UTMASTG/UBLKCP, commit -> UTCBAR), from the shipped library.
usage: sass_opcodes.py [library.so] > profiles/sass_opcodes.txt"""
import collections
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = sys.argv[1] if len(sys.argv) > 1 else os.path.join(ROOT, "deepgemm_ascend_b200", "lib", "libdga_b200.so")
WATCH = ("UTCHMMA", "UTCQMMA", "UTCOMMA", "UTCMXQMMA", "LDTM", "STTM", "UTCCP", "UTCBAR", "UTMALDG", "UTMASTG", "UTMAPF", "UBLKCP",
         "SYNCS", "FFMA2", "HMMA", "HGMMA")


def main():
    sass = subprocess.run(["cuobjdump", "-sass", LIB], capture_output=True, text=True, check=True).stdout
    per = collections.OrderedDict()
    name = None
    for line in sass.splitlines():
        m = re.match(r"\s*Function : (\S+)", line)
        if m:
            name = subprocess.run(["c++filt", m.group(1)], capture_output=True, text=True).stdout.strip()
            name = re.sub(r"\(CUtensorMap_st.*", "", name)
            per[name] = collections.Counter()
            continue
        m = re.match(r"\s*/\*[0-9a-f]+\*/\s+(?:@!?U?P\d+\s+)?([A-Z0-9_]+)", line)
        if m and name is not None:
            op = m.group(1)
            for w in WATCH:
                if op.startswith(w):
                    per[name][w] += 1
    total = collections.Counter()
    print("# SASS mnemonic counts per kernel of %s" % os.path.relpath(LIB, ROOT))
    print("# (cuobjdump -sass; legacy tensor paths would show HMMA / HGMMA: none)")
    for kname, cnt in per.items():
        if not cnt:
            continue
        total.update(cnt)
        print("%s\n    %s" % (kname, "  ".join("%s %d" % (w, cnt[w]) for w in WATCH if cnt[w])))
    print("TOTAL\n    %s" % "  ".join("%s %d" % (w, total[w]) for w in WATCH if total[w]))


if __name__ == "__main__":
    main()
