"""Summarise a CPOLY_TC_TRACE event log (stderr of a run with CPOLY_TC_TRACE=<launch index>): per-role step timing of CTA 0.
Debugging aid for the pipeline of csrc/dcn_tc_win.cu.   python tools/trace_summary.py gpurun_out/trace.txt"""
import collections
import sys


def main():
    ev = collections.defaultdict(list)
    for line in open(sys.argv[1]):
        if line.startswith("TRACE"):
            _, r, tag, t = line.split()
            ev[int(r)].append((int(tag), int(t)))
    for r in sorted(ev):
        print("role", r, "events", len(ev[r]), "first", ev[r][0], "last", ev[r][-1])
    s = ev[0]
    t1 = [t for tag, t in s if tag == 1]
    t2 = [t for tag, t in s if tag == 2]
    print("sampler step period     :", [b - a for a, b in zip(t1, t1[1:])])
    print("sampler blend..arrive   :", [b - a for a, b in zip(t1, t2)])
    print("sampler arrive..emptyA  :", [b - a for a, b in zip(t2, t1[1:])])
    m = ev[2]
    d = collections.defaultdict(list)
    for (ta, a), (tb, b) in zip(m, m[1:]):
        d[(ta, tb)].append(b - a)
    for k in sorted(d):
        print("mma", k, d[k][:24])
    print("loader", ev[1][:40])
    print("epilogue", ev[3])


if __name__ == "__main__":
    main()
