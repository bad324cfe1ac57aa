"""Per-kernel SASS evidence for the sm_100a library: counts of the tcgen05 / TMEM / TMA / PDL mnemonics in every kernel of
libubresnet_b200.so (cuobjdump -sass).  usage: python tools/sass_summary.py > profiles/r02_sass_summary.txt
Mnemonics (B200_PROFILING.md): UTCHMMA = tcgen05.mma, LDTM / STTM = tcgen05.ld / st (TMEM), UTMALDG / UTMASTG = TMA load /
store (cp.async.bulk.tensor), UTCBAR = tcgen05.commit, SYNCS = mbarrier ops, PREEXIT / ACQBULK = griddepcontrol.launch_dependents / .wait (PDL)."""
import collections
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = os.path.join(ROOT, "ubresnet_b200", "lib", "libubresnet_b200.so")
KEYS = ["UTCHMMA", "LDTM", "STTM", "UTMALDG", "UTMASTG", "UTCBAR", "SYNCS", "PREEXIT", "ACQBULK", "HMMA", "FFMA", "LDG", "STG", "SHFL", "ATOM"]


def main():
    out = subprocess.run(["cuobjdump", "-sass", LIB], capture_output=True, text=True, check=True).stdout
    demangle = {}
    counts = collections.OrderedDict()
    cur = None
    for line in out.splitlines():
        m = re.match(r"\s*Function : (\S+)", line)
        if m:
            cur = m.group(1)
            counts[cur] = collections.Counter()
            continue
        if cur is None:
            continue
        m = re.match(r"\s*/\*[0-9a-f]+\*/\s+(?:@!?U?P\d+\s+)?([A-Z0-9_.]+)", line)
        if m:
            op = m.group(1)
            counts[cur]["_total"] += 1
            base = op.split(".")[0]
            for k in KEYS:
                if base == k or (k == "UTCHMMA" and base.startswith("UTCHMMA")) or (k == "ATOM" and base.startswith("ATOM")):
                    counts[cur][k] += 1
    names = list(counts)
    try:
        dem = subprocess.run(["c++filt"] + names, capture_output=True, text=True).stdout.splitlines()
        demangle = dict(zip(names, dem))
    except Exception:
        pass
    print("# SASS summary of %s (sm_100a), %d kernels" % (os.path.relpath(LIB, ROOT), len(names)))
    print("# columns: total instructions | " + " ".join(KEYS))
    tot = collections.Counter()
    for n in names:
        c = counts[n]
        short = re.sub(r"\(.*", "", demangle.get(n, n)).replace("void ", "").replace("ubr::", "")
        print("%-52s %6d | %s" % (short[:52], c["_total"], " ".join("%s=%d" % (k, c[k]) for k in KEYS if c[k])))
        tot.update(c)
    print("# library totals: " + " ".join("%s=%d" % (k, tot[k]) for k in KEYS))


if __name__ == "__main__":
    main()
