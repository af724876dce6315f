"""Per-kernel SASS opcode histogram of lib/libaphrodite_b200.so (cuobjdump -sass): how often each kernel family uses the
tcgen05 tensor core (UTCHMMA), tensor memory (STTM / LDTM), TMA (UTMALDG tiled, UBLKCP bulk), mma.sync (HMMA), ldmatrix
(LDSM), cluster barriers (UCGABAR_ARV / _WAIT) and PDL (ACQBULK = griddepcontrol.wait, PREEXIT = launch_dependents). Evidence
that the paths named in DESIGN.md are what the compiler emitted.   python tools/sass_histogram.py > profiles/<name>.txt"""
import collections
import os
import re
import subprocess
import sys

LIB = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "aphrodite-fork_b200", "lib", "libaphrodite_b200.so")
OPS = ["UTCHMMA", "UTCQMMA", "STTM", "LDTM", "UTMALDG", "UBLKCP", "UTMAPF", "HMMA", "LDSM", "MOVM", "SYNCS", "UCGABAR_ARV", "UCGABAR_WAIT",
       "ACQBULK", "PREEXIT", "UTCBAR", "ELECT", "REDUX", "ATOM", "RED", "F2FP", "MUFU.EX2", "LDGMC", "STGMC", "NANOSLEEP"]


def main():
    out = subprocess.run(["cuobjdump", "-sass", LIB], capture_output=True, text=True, check=True).stdout
    per = collections.OrderedDict()
    name = None
    for line in out.splitlines():
        m = re.match(r"\s*Function : (\S+)", line)
        if m:
            name = subprocess.run(["c++filt", m.group(1)], capture_output=True, text=True).stdout.strip()
            name = name.replace("(anonymous namespace)::", "")
            name = re.sub(r"\(.*", "", name)
            name = re.sub(r"^void\s+", "", name)
            per[name] = collections.Counter()
            continue
        if name is None:
            continue
        m = re.search(r"^\s*/\*[0-9a-f]{4}\*/\s+(?:@!?U?P\d+\s+)?([A-Z0-9_.]+)", line)
        if not m:
            continue
        op = m.group(1)
        per[name]["_total"] += 1
        for o in OPS:
            if op == o or op.startswith(o + "."):
                per[name][o] += 1
    # one line per kernel family (template arguments collapsed), counts summed over its instantiations
    fam = collections.OrderedDict()
    for n, c in per.items():
        f = re.sub(r"<.*", "", n)
        fam.setdefault(f, [0, collections.Counter()])
        fam[f][0] += 1
        fam[f][1].update(c)
    print(f"# {os.path.relpath(LIB)}: SASS opcode counts per kernel family (instantiations summed); cuobjdump -sass, sm_100a")
    tot = collections.Counter()
    for f, (n, c) in fam.items():
        tot.update(c)
        items = " ".join(f"{o}={c[o]}" for o in OPS if c[o])
        print(f"{f}  [{n} instantiation(s), {c['_total']} instructions]  {items}")
    print("# whole library: " + " ".join(f"{o}={tot[o]}" for o in OPS if tot[o]))


if __name__ == "__main__":
    main()
