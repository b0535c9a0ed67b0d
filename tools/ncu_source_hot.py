"""Hot source lines of an ncu report: per kernel, the source lines with most stall samples / instructions.
    ncu -i X.ncu-rep --page source --csv --print-source cuda,sass > X_src.csv ; python tools/ncu_source_hot.py X_src.csv [N]"""
import csv
import sys
import collections


def main():
    n = int(sys.argv[2]) if len(sys.argv) > 2 else 25
    fn = path = None
    hdr = None
    agg = collections.defaultdict(lambda: collections.defaultdict(lambda: [0, 0, ""]))   # fn -> (file,line) -> [samples, inst, text]
    for r in csv.reader(open(sys.argv[1], errors="replace")):
        if not r:
            continue
        if r[0] == "File Path":
            path = r[1].split("/")[-1]; continue
        if r[0] == "Function Name":
            fn = r[1][:60]; continue
        if r[0] == "Line No":
            hdr = r; continue
        if hdr is None or r[0] == "":
            continue
        try:
            samples = int(r[hdr.index("# Samples")]); inst = int(r[hdr.index("Instructions Executed")])
        except ValueError:
            continue
        a = agg[fn][(path, r[0])]
        a[0] += samples; a[1] += inst; a[2] = r[1][:90]
    for fn, lines in agg.items():
        ts = sum(v[0] for v in lines.values()); ti = sum(v[1] for v in lines.values())
        print(f"=== {fn}  samples={ts} warp-inst={ti}")
        for (p, ln), v in sorted(lines.items(), key=lambda kv: -kv[1][0])[:n]:
            print(f"  {100*v[0]/max(ts,1):5.1f}% smp {100*v[1]/max(ti,1):5.1f}% inst  {p}:{ln}  {v[2]}")


if __name__ == "__main__":
    main()
