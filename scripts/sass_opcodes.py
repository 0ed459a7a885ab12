"""Per-kernel SASS opcode evidence for profiles/: counts of the mnemonics that prove tcgen05 / TMEM / TMA use
(UTCHMMA, LDTM, STTM, UTMALDG, UTMAPF, UTCBAR, SYNCS) and of the legacy tensor path (HMMA), per kernel of
libjanus_b200.so.

    python scripts/sass_opcodes.py > profiles/r02_sass_opcodes.txt
"""
import collections
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = os.path.join(ROOT, "janus_b200", "_lib", "libjanus_b200.so")
WATCH = ("UTCHMMA", "UTCQMMA", "LDTM", "STTM", "UTMALDG", "UTMASTG", "UTMAPF", "UTCBAR", "SYNCS", "HMMA", "FFMA", "FFMA2",
         "FADD", "FMUL", "MUFU", "LDS", "STS", "LDG", "STG")


def main():
    out = subprocess.run(["cuobjdump", "-sass", LIB], capture_output=True, text=True).stdout
    fns, cur = collections.OrderedDict(), None
    for ln in out.splitlines():
        m = re.match(r"\s*Function : (\S+)", ln)
        if m:
            cur = m.group(1)
            fns[cur] = collections.Counter()
            continue
        if cur:
            m = re.match(r"\s*/\*[0-9a-f]{4}\*/\s+(?:@!?U?P\d+\s+)?([A-Z][A-Z0-9_]*)", ln)
            if m:
                fns[cur][m.group(1)] += 1
                fns[cur]["_total"] += 1
    demangle = subprocess.run(["c++filt"], input="\n".join(fns), capture_output=True, text=True).stdout.splitlines()
    print(f"# cuobjdump -sass {os.path.relpath(LIB, ROOT)} (sm_100a): opcode counts per kernel")
    print("# " + " ".join(f"{w:>8s}" for w in ("total",) + WATCH) + "  kernel")
    for (name, c), dn in zip(fns.items(), demangle):
        dn = re.sub(r"\(.*", "", dn)
        print("  " + " ".join(f"{c.get(w if w != 'total' else '_total', 0):8d}" for w in ("total",) + WATCH) + "  " + dn[:110])


if __name__ == "__main__":
    sys.exit(main())
