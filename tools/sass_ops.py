"""Per-kernel SASS opcode histogram of libflyb200.so -> profiles/sass_ops.txt (`python tools/sass_ops.py`; CPU only).

Proves which kernels carry tcgen05 / TMEM / bulk-copy instructions without anyone having to disassemble the library:
UTCHMMA = tcgen05.mma, LDTM / STTM = tcgen05.ld / st, UTCBAR = tcgen05.commit, UBLKCP = cp.async.bulk (1-D TMA),
UTMALDG / UTMASTG = tensor-map TMA, SYNCS = mbarrier operations."""
import collections
import os
import re
import subprocess

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = os.path.join(ROOT, "fly_bproject_b200", "libflyb200.so")
OPS = ["UTCHMMA", "UTCQMMA", "UTCBAR", "LDTM", "STTM", "UBLKCP", "UTMALDG", "UTMASTG", "SYNCS", "HMMA", "FFMA", "LDG",
       "STG", "LDS", "STS", "ATOMG", "REDG", "BAR", "MUFU"]


def main():
    out = subprocess.run(["cuobjdump", "-sass", LIB], capture_output=True, text=True, check=True).stdout
    cur, hist = None, collections.OrderedDict()
    for ln in out.splitlines():
        m = re.search(r"Function : (\S+)", ln)
        if m:
            cur = m.group(1)
            hist[cur] = collections.Counter()
            continue
        m = re.search(r"^\s+/\*[0-9a-f]+\*/\s+(?:@!?U?P\d+\s+)?([A-Z0-9_]+)", ln) if cur else None
        if m:
            hist[cur]["_total"] += 1
            if m.group(1) in OPS:
                hist[cur][m.group(1)] += 1
    names = subprocess.run(["c++filt", "-p"], input="\n".join(hist), capture_output=True, text=True).stdout.split("\n")
    with open(os.path.join(ROOT, "profiles", "sass_ops.txt"), "w") as f:
        f.write("# SASS opcode histogram per kernel of fly_bproject_b200/libflyb200.so (sm_100a): cuobjdump -sass, tools/sass_ops.py\n")
        f.write("# UTCHMMA = tcgen05.mma, LDTM/STTM = tcgen05.ld/st (TMEM), UTCBAR = tcgen05.commit, UBLKCP = cp.async.bulk (1-D TMA),\n"
                "# UTMALDG/UTMASTG = tensor-map TMA (none: 1-D bulk copies only), SYNCS = mbarrier ops\n")
        f.write("%-72s %7s " % ("kernel", "instrs") + " ".join("%7s" % o for o in OPS) + "\n")
        tot = collections.Counter()
        for (k, h), name in zip(hist.items(), names):
            f.write("%-72s %7d " % (name[:72], h["_total"]) + " ".join("%7d" % h[o] for o in OPS) + "\n")
            tot.update(h)
        f.write("%-72s %7d " % ("TOTAL", tot["_total"]) + " ".join("%7d" % tot[o] for o in OPS) + "\n")


if __name__ == "__main__":
    main()
