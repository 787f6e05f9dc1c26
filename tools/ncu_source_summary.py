#!/usr/bin/env python
"""Summarise `ncu --page source --csv --print-source cuda,sass` output: top source lines by stall samples.

  ncu -i rep.ncu-rep --page source --csv --print-source cuda,sass --kernel-name regex:X --launch-skip N --launch-count 1 > s.csv
  python tools/ncu_source_summary.py s.csv [topN]
"""
import csv
import sys


def main():
    path = sys.argv[1]
    top = int(sys.argv[2]) if len(sys.argv) > 2 else 40
    cur_file = None
    hdr = None
    out = []
    for r in csv.reader(open(path, errors="replace")):
        if not r:
            continue
        if r[0] == "File Path":
            cur_file = r[1].split("/")[-1]
            continue
        if r[0] == "Line No":
            hdr = r
            continue
        if hdr is None or r[0] in ("Function Name",) or not r[0].isdigit():
            continue
        d = dict(zip(hdr, r))  # duplicate 'Source' key: second (SASS) wins, so re-read line text
        line_txt = r[1]
        try:
            samples = int(r[hdr.index("# Samples")])
        except ValueError:
            continue
        stalls = {k[6:]: int(v) for k, v in d.items() if k.startswith("stall_") and "Not Issued" not in k and v.isdigit() and int(v) > 0}
        inst = d.get("Instructions Executed", "0")
        out.append((samples, cur_file, r[0], line_txt.strip()[:110], inst, stalls))
    tot = sum(o[0] for o in out)
    print(f"total samples {tot}")
    for s, f, ln, txt, inst, st in sorted(out, key=lambda o: -o[0])[:top]:
        tops = ", ".join(f"{k}={v}" for k, v in sorted(st.items(), key=lambda kv: -kv[1])[:4])
        print(f"{100*s/max(tot,1):5.1f}% {s:7d} {f}:{ln:>4} inst={inst:>9} | {txt}\n          {tops}")


if __name__ == "__main__":
    main()
