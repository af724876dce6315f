"""Summarise `ncu -i X.ncu-rep --page source --csv` (stdin or file): warp-stall samples per SASS opcode with their top
stall reasons, and the hottest instructions.   ncu -i prof.ncu-rep --page source --csv | python tools/ncu_source_summary.py"""
import collections
import csv
import re
import sys


def main():
    rows = list(csv.reader(open(sys.argv[1]) if len(sys.argv) > 1 else sys.stdin))
    hi = [i for i, r in enumerate(rows) if r and r[0] == "Address"][0]
    hdr = rows[hi]
    ix = {h: i for i, h in enumerate(hdr)}
    stalls = [h for h in hdr if h.startswith("stall_") and "Not Issued" not in h]
    tot, execs, by_op, lines, alls = collections.Counter(), collections.Counter(), collections.defaultdict(collections.Counter), [], 0

    def num(v):
        try:
            return int(float(v))
        except ValueError:
            return 0

    for r in rows[hi + 1:]:
        if len(r) < len(hdr) or r[0] == "Address":
            continue
        src = r[ix["Source"]]
        m = re.match(r"\s*(@!?U?P\d+\s+)?([A-Z0-9_.]+)", src)
        if not m:
            continue
        op = ".".join(m.group(2).split(".")[:2]) if m.group(2).startswith(("LDS", "HMMA", "SYNCS", "UTC", "LDSM")) else m.group(2).split(".")[0]
        n = num(r[ix["# Samples"]])
        execs[op] += num(r[ix["Instructions Executed"]])
        tot[op] += n
        alls += n
        st = {s: num(r[ix[s]]) for s in stalls}
        for s, v in st.items():
            if v:
                by_op[op][s] += v
        lines.append((n, src[:100], st))
    print(f"{rows[0][1][:120] if len(rows[0]) > 1 else ''}\ntotal stall samples {alls}")
    for op, n in tot.most_common(int(sys.argv[2]) if len(sys.argv) > 2 else 16):
        top = ", ".join(f"{k[6:]}={v}" for k, v in by_op[op].most_common(4))
        print(f"{op:14s} samples {n:6d} ({100 * n / max(alls, 1):4.1f}%)  executed {execs[op]:9d}   {top}")
    print("hottest instructions:")
    for n, src, st in sorted(lines, key=lambda x: -x[0])[:12]:
        print(f"  {n:5d}  {src.strip():80s} " + ", ".join(f"{k[6:]}={v}" for k, v in sorted(st.items(), key=lambda kv: -kv[1])[:3] if v))


if __name__ == "__main__":
    main()
