import numpy as np, random, math, collections
from pil_rules import *
from pil_fill_restatement import RU, RD
f32=np.float32
W,H=120,100
def runs(row):
    xs=np.nonzero(row)[0]
    return (int(xs[0]), int(xs[-1])) if len(xs) else None
random.seed(2)
# triangles: V is the unique top (or bottom) vertex; edges e_first (earlier in edge table) and e_second
stats=collections.defaultdict(lambda: collections.Counter())
tot=collections.Counter()
for t in range(6000):
    top = random.random()<0.5
    V=(random.randint(30,90), random.randint(30,70))
    sgn = random.choice([1,-1])     # side where the polygon lies
    def other():
        dy=random.randint(1,25)*(1 if top else -1)
        dxs=random.randint(0,50)*sgn
        return (V[0]+dxs, V[1]+dy)
    A,B=other(),other()
    if A[1]==B[1] and A[0]==B[0]: continue
    order=random.choice([0,1,2])
    P=[V,A,B]; P=P[order:]+P[:order]
    if random.random()<0.5: P=P[::-1]
    pf=pil_fill_only(P,W,H)
    r=runs(pf[V[1]])
    if r is None: continue
    # edge table order
    n=3; E=[(P[i],P[(i+1)%3]) for i in range(3)]
    inc=[(i,e) for i,e in enumerate(E) if V in e and e[0][1]!=e[1][1]]
    if len(inc)!=2: continue
    (i1,e1),(i2,e2)=inc   # e1 earlier in table = "other", e2 = "current"
    def dxo(e): return f32(f32(e[1][0]-e[0][0])/f32(e[1][1]-e[0][1]))
    def ex(e,y): return f32(f32(y-e[0][1])*dxo(e)+f32(e[0][0]))
    ya=V[1]+(1 if top else -1)
    # adjacent row must be inside both edges
    if not all(min(e[0][1],e[1][1])<=ya<=max(e[0][1],e[1][1]) for e in (e1,e2)): continue
    cur=ex(e2,ya); oth=ex(e1,ya)
    dcur=dxo(e2); doth=dxo(e1)
    if (dcur>0)!=(doth>0) or dcur==0 or doth==0: continue
    key=('top' if top else 'bottom', 'dx>0' if dcur>0 else 'dx<0')
    ext = r[1] if r[0]==V[0] else r[0]      # the non-vertex end of the span on V's row
    if r[0]!=V[0] and r[1]!=V[0]: continue
    side = 'right' if r[0]==V[0] else 'left'
    cands={'none':V[0],
      'RD(min-1)':RD(min(cur,oth)-1),'RD(min)':RD(min(cur,oth)),'RD(cur-1)':RD(cur-1),'RD(oth-1)':RD(oth-1),'RU(min)-1':RU(min(cur,oth))-1,
      'RU(max+1)':RU(max(cur,oth)+1),'RU(max)':RU(max(cur,oth)),'RU(cur+1)':RU(cur+1),'RU(oth+1)':RU(oth+1),'RD(max)+1':RD(max(cur,oth))+1}
    tot[key]+=1
    for k,v in cands.items():
        # candidate predicts span end: if candidate on the far side of V use it else V
        if k.startswith('RD') or k.startswith('RU(min'):   # right-side extension candidates
            pred = (V[0], max(V[0], v))
        else:
            pred = (min(V[0], v), V[0])
        if k=='none': pred=(V[0],V[0])
        stats[key][k]+= (pred==r)
for key in sorted(tot):
    print(key, tot[key], {k:v for k,v in stats[key].most_common(6)})
