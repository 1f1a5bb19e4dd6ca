"""Per-source-line instruction counts and stall samples from an ncu report (first profiled launch only).

    python tools/ncu_lines.py gpurun_out/prof.ncu-rep [top_n] [kernel-name regex]

Needs `ncu` on PATH (no GPU); the kernels must have been built with -lineinfo."""
import collections
import csv
import io
import os
import subprocess
import sys


def main():
    rep = sys.argv[1]
    top = int(sys.argv[2]) if len(sys.argv) > 2 else 40
    pick = ["-k", "regex:" + sys.argv[3]] if len(sys.argv) > 3 else []
    out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "cuda,sass", *pick],
                         capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(out)))
    data = collections.OrderedDict()
    fname, hdr, seen_files = None, None, set()
    for r in rows:
        if len(r) == 2 and r[0] == "File Path":
            if r[1] in seen_files:      # second launch starts: stop
                break
            seen_files.add(r[1])
            fname = os.path.basename(r[1])
            continue
        if r and r[0] == "Line No":
            hdr = r
            i_inst, i_samp = hdr.index("Instructions Executed"), hdr.index("# Samples")
            continue
        if hdr is None or len(r) < len(hdr) or not r[0].isdigit():
            continue
        try:
            inst, samp = int(r[i_inst]), int(r[i_samp])
        except ValueError:
            continue
        key = (fname, int(r[0]))
        e = data.setdefault(key, [0, 0, r[1].strip()])
        e[0] += inst
        e[1] += samp
    tot = sum(v[0] for v in data.values()) or 1
    ts = sum(v[1] for v in data.values()) or 1
    print(f"total warp-instructions {tot}, stall samples {ts}")
    for (f, l), (i, s, src) in sorted(data.items(), key=lambda kv: -kv[1][0])[:top]:
        print(f"{i:10d} {100 * i / tot:5.1f}%  samp {100 * s / ts:5.1f}%  {f}:{l}  {src[:100]}")


if __name__ == "__main__":
    main()
