"""Where a kernel's instructions are executed: `ncu -i X.ncu-rep --page source --csv` (file or stdin, first kernel of
the file) cut into runs of 64 SASS instructions, each with its share of the executed warp instructions, its stall
samples and the landmark opcodes inside (TMA, tcgen05, barriers) that tell which warp role the run belongs to.
  ncu -i prof.ncu-rep --page source --csv | python tools/ncu_region_breakdown.py [min_share_percent]"""
import csv
import sys

LANDMARKS = ("STTM", "LDTM", "UTCHMMA", "UTCBAR", "UTMALDG", "UBLKCP", "SYNCS.PHASECHK", "SYNCS.ARRIVE", "BAR.SYNC", "NANOSLEEP",
             "HMMA", "LDSM", "EXIT", "ATOM", "RED", "STG", "LDG", "UCGABAR")


def main():
    args = [a for a in sys.argv[1:]]
    src = open(args[0]) if args and not args[0].replace(".", "").isdigit() else sys.stdin
    thr = float(args[-1]) if args and args[-1].replace(".", "").isdigit() else 0.3
    rows, hdr, name = [], None, ""
    for r in csv.reader(src):
        if r and r[0] == "Kernel Name":
            if rows:
                break  # first kernel only
            name = r[1]
        elif r and r[0] == "Address":
            hdr = r
        elif hdr and len(r) >= len(hdr) - 2:
            rows.append(r)
    ie, ism, isrc, ia = hdr.index("Instructions Executed"), hdr.index("# Samples"), hdr.index("Source"), hdr.index("Address")
    num = lambda v: int(float(v)) if v else 0
    tot = sum(num(r[ie]) for r in rows)
    print(name[:140])
    print(f"SASS instructions {len(rows)}, executed warp instructions {tot}")
    for b in range(0, len(rows), 64):
        chunk = rows[b:b + 64]
        e, sm = sum(num(r[ie]) for r in chunk), sum(num(r[ism]) for r in chunk)
        if e < tot * thr / 100:
            continue
        marks = sorted({k for r in chunk for k in LANDMARKS if k in r[isrc]})
        print(f"{b:6d} @{chunk[0][ia][-5:]}  executed {e:10d} ({100 * e / tot:5.1f} %)  samples {sm:5d}  {', '.join(marks)}")


if __name__ == "__main__":
    main()
