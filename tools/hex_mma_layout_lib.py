"""Bank-conflict model of hex_mma_kernel's ring layout (fem_b200/csrc/hex_mma.cu): the symbolic shared-memory accesses of
the gather warps (`runs`) for both node-row parities.  tools/hex_mma_anneal.py anneals the 36-entry permutation
`c_perm`, the cell pitch and the ring-slot pitch against them plus the MMA warps' stores."""
def sym(i, j): return {(0, 0): 0, (0, 1): 1, (0, 2): 2, (1, 1): 3, (1, 2): 4, (2, 2): 5}[(i, j)]
def lane_value(l):
    if l < 9: return (l // 3 - 1, 0, l % 3)
    if l < 27:
        k = l - 9
        return ((k // 3) % 3 - 1, -1 if k < 9 else 1, k % 3)
    return None
import random, sys

def raw(P,m,n):
    s=n&1; q=n>>1
    assert not ((m&1)==P and s==P) and not (m>=4 and q>=2)
    if s!=P: return q*8+m if q<2 else 16+(q-2)*4+m
    r2=m>>1
    return 24+q*4+r2 if q<2 else 32+(q-2)*2+r2
# symbolic accesses
runs=[]  # each run: list of (lane, slot, ox, t, P, rawidx)
for par in range(2):
    for g in range(3):
        bzr,bzc=[(1,0),(0,0),(0,1)][g]
        for i in range(3):
            for c in range(2):
                acc=[]
                for l in range(27):
                    dy,dx,j=lane_value(l)
                    cons=[ox for ox in range(2) if 0<=(1-ox)+dx<=1]
                    if c>=len(cons): continue
                    ox=cons[c]; bxr=1-ox; bxc=bxr+dx
                    if dy==-1: slot=0; P=par; pc=par^1
                    else: slot=1; P=par^1; pc=par^(dy&1)
                    m=4*bzr+2*bxr+par; n=4*bzc+2*bxc+pc
                    if i<=j: t=sym(i,j); r=raw(P,m,n)
                    else: t=sym(j,i); r=raw(P,n,m)
                    acc.append((l,slot,ox,t,P,r))
                runs.append(acc)
def cost(perm,TP,CP,SP):
    tot=0
    for acc in runs:
        for h in range(2):
            banks={}
            for (l,slot,ox,t,P,r) in acc:
                if (l>>4)==h:
                    a=slot*SP+ox*CP+t*TP+perm[P][r]
                    banks.setdefault(a&15,set()).add(a)
            if banks: tot+=max(len(s) for s in banks.values())
    return tot/2
