"""Exhaustive search over one- and two-field XOR swizzles of the FFT exchange buffer (fft_exact.cuh Plan::phys):
models every store / load pattern of every pass for a warp (16-lane phases, 64-bit accesses, the kernels' window pitch)
and prints the current and the best (worst degree, summed degree) per SF.  CPU only."""
import itertools, sys
def plan(SF):
    N=1<<SF; R=16; T=N//R; NP=(SF+3)//4; W0=SF-4*(NP-1); STAGES=(SF+1)//2; ODD=SF&1
    def pos_from_input(i):
        o=0
        for j in range(STAGES):
            last2 = ODD and j==STAGES-1
            q = (i>>(2*j))&1 if last2 else (i>>(2*j))&3
            shift = 0 if last2 else SF-2*(j+1)
            o |= q<<shift
        return o
    def pos_in_pass(p,tid,e):
        lo=W0+4*(p-1)
        return ((tid>>lo)<<(lo+4))|(e<<lo)|(tid&((1<<lo)-1))
    pats=[]   # list of lists: addresses for all threads of one window, per (exchange, r)
    for p in range(1,NP):
        for r in range(R):
            pats.append([ (pos_from_input(t+T*r) if p==1 else pos_in_pass(p-1,t,r)) for t in range(T)])
        for r in range(R):
            pats.append([pos_in_pass(p,t,r) for t in range(T)])
    return N,T,pats
def worst(SF, phys, pitch_extra=8):
    N,T,pats=plan(SF)
    G=max(1,32//T)             # windows sharing a warp
    PITCH=N+pitch_extra
    w=1; total=0
    for pat in pats:
        # lanes of a warp: for T>=32 consecutive tids; for T<32, G groups side by side
        if T>=32:
            warps=[[phys(pat[t]) for t in range(w0,w0+32)] for w0 in range(0,T,32)]
        else:
            warps=[[g*PITCH+phys(pat[t]) for g in range(G) for t in range(T)]]
        for lanes in warps:
            for h in (lanes[:16],lanes[16:]):
                banks={}
                for a in h: banks.setdefault(a&15,set()).add(a)
                m=max(len(v) for v in banks.values())
                w=max(w,m); total+=m
    return w,total
def cur(SF):
    if SF>=8: return lambda o:o^((o>>(SF-4))&15)
    if SF==7: return lambda o:o^((o>>4)&7)
    return lambda o:o
for SF in range(7,13):
    print(SF,'current',worst(SF,cur(SF)))
    best=None
    N=1<<SF
    # candidates: XOR of up to two bit-fields (src >= 4) into low 4 bits
    fields=[(s,wd,d) for s in range(4,SF) for wd in range(1,5) if s+wd<=SF for d in range(0,5-wd)]
    cands=[()]+[(f,) for f in fields]+[(f,g) for f in fields for g in fields if f<g]
    for c in cands:
        def phys(o,c=c):
            x=o
            for (s,wd,d) in c: x^=((o>>s)&((1<<wd)-1))<<d
            return x
        r=worst(SF,phys)
        if best is None or r<best[0]: best=(r,c)
    print(SF,'best',best)
