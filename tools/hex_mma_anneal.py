import random, sys
from hex_mma_layout_lib import runs, raw
# emit store groups: per P, per slot-kind: list of (lane, rawidx)
stores=[]
for P in range(2):
    for kind in range(2):   # 0: slot 1-P (col -y), 1: slot P (col +y)
        acc=[]
        for l in range(32):
            m=l>>2; q=l&3
            s=(1-P) if kind==0 else P
            n=2*q+s
            if ((m&1)==P and s==P) or (m>=4 and q>=2): continue
            acc.append((l,P,raw(P,m,n)))
        stores.append(acc)
def wf(pairs):
    tot=0
    for h in range(2):
        banks={}
        for l,a in pairs:
            if (l>>4)==h: banks.setdefault(a&15,set()).add(a)
        if banks: tot+=max(len(s) for s in banks.values())
    return tot
def cost(perm,TP,CP,SP):
    g=0
    for acc in runs:
        g+=wf([(l,slot*SP+ox*CP+t*TP+perm[P][r]) for (l,slot,ox,t,P,r) in acc])
    s=0
    for acc in stores:
        s+=wf([(l,perm[P][r]) for (l,P,r) in acc])   # same for each tile: x6 tiles, /2 parities
    return g/2 + s*6/2, g/2, s*6/2
seed=int(sys.argv[1]); random.seed(seed)
TP=int(sys.argv[2])
best=None
for trial in range(3):
    CP=6*TP+random.randrange(0,10); SP=8*CP+random.randrange(0,16)
    perm=[list(range(TP)),list(range(TP))]
    cur=cost(perm,TP,CP,SP); T=1.5
    for it in range(12000):
        kind=random.random()
        if kind<0.85:
            P=random.randrange(2); a=random.randrange(TP); b=random.randrange(TP)
            perm[P][a],perm[P][b]=perm[P][b],perm[P][a]
            c2=cost(perm,TP,CP,SP)
            if c2[0]<=cur[0] or random.random()<pow(2.718,-(c2[0]-cur[0])/T): cur=c2
            else: perm[P][a],perm[P][b]=perm[P][b],perm[P][a]
        else:
            o=(CP,SP)
            if kind<0.92: CP=6*TP+random.randrange(0,12)
            else: SP=8*CP+random.randrange(0,16)
            c2=cost(perm,TP,CP,SP)
            if c2[0]<=cur[0]: cur=c2
            else: CP,SP=o
        T=max(0.03,T*0.9995)
    print(seed,TP,CP,SP,cur,flush=True)
    if best is None or cur[0]<best[0][0]: best=(cur,TP,CP,SP,[p[:] for p in perm])
print("BEST",best)
