"""Summarise a CPOLY_TC_TRACE log of csrc/dcn_tc.cu (CTA 0): python tools/trace_gather.py file"""
import collections, sys
ev = collections.defaultdict(list)
for line in open(sys.argv[1]):
    if line.startswith("TRACE"):
        _, r, tag, t = line.split()
        ev[int(r)].append((int(tag), int(t)))
for r in sorted(ev):
    print("role", r, "n", len(ev[r]), "first", ev[r][0], "last", ev[r][-1])
def seq(r, tag): return [t for g, t in ev[r] if g == tag]
for r in (0, 4):
    a, b, c = seq(r, 1), seq(r, 2), seq(r, 3)
    print(f"sampler{r} period      ", [y - x for x, y in zip(a, a[1:])])
    print(f"sampler{r} wait->arrive", [y - x for x, y in zip(a, b)])
    print(f"sampler{r} arrive->next", [y - x for x, y in zip(b, c)])
    print(f"sampler{r} next->wait ok", [y - x for x, y in zip(c, a[1:])])
m1, m2, m3 = seq(2, 1), seq(2, 2), seq(2, 3)
print("mma period  ", [y - x for x, y in zip(m3, m3[1:])])
print("mma B wait  ", [y - x for x, y in zip(m3, m1[1:])])
print("mma A wait  ", [y - x for x, y in zip(m1, m2)])
print("mma issue   ", [y - x for x, y in zip(m2, m3)])
l = seq(1, 1)
print("loader period", [y - x for x, y in zip(l, l[1:])])
print("param", ev[3])
