import sys, collections
sys.path.insert(0,'/root/repo/tools/scratch')
from sched_sim import parse, simulate
srcs,kinds,outs=parse(sys.argv[1])
K=int(sys.argv[2]) if len(sys.argv)>2 else 4
n=len(srcs); no=len(outs)
# blocks: CSR outputs 0..909 split evenly by index; the rest go with the last block
NCSR=910
def blk(oi): return min(K-1, oi*K//NCSR) if oi<NCSR else K-1
mask=[0]*n
for oi,o in enumerate(outs): mask[o]|=1<<blk(oi)
for i in range(n-1,-1,-1):
    for s in srcs[i]: mask[s]|=mask[i]
users=[[] for _ in range(n)]
for i in range(n):
    for s in srcs[i]: users[s].append(i)
pc=lambda m: bin(m).count("1")
shared=[i for i in range(n) if pc(mask[i])>=2 and kinds[i] not in("ld","const")]
print("K",K,"nodes",n,"shared compute nodes",len(shared), "by popcount", collections.Counter(pc(mask[i]) for i in shared))
# frontier: shared nodes with a user whose mask is a strict subset / or that are outputs
front=[i for i in shared if any(mask[u]!=mask[i] for u in users[i])]
print("frontier (shared values consumed by a node with a smaller block set):",len(front))
# frontier to *private* (popcount-1) users
front1=[i for i in shared if any(pc(mask[u])==1 for u in users[i])]
print("shared values consumed directly by private nodes:",len(front1))
# per block: private ops count and max live if shared values are free
for b in range(K):
    bouts=[o for oi,o in enumerate(outs) if blk(oi)==b]
    # treat shared nodes as loads
    k2=list(kinds)
    s2=[list(x) for x in srcs]
    for i in shared: k2[i]="ld"; s2[i]=[]
    mx,prof,nops=simulate(s2,k2,bouts)
    priv=sum(1 for i in range(n) if mask[i]==(1<<b) and kinds[i] not in("ld","const"))
    need=len(set(i for i in shared if any(mask[u]==(1<<b) for u in users[i])))
    print(" block",b,"outs",len(bouts),"private ops",priv,"max live private",mx,"shared values read",need)
# full recompute cost per block (what pieces do now)
for b in range(K):
    tot=sum(1 for i in range(n) if mask[i]>>b&1 and kinds[i] not in("ld","const"))
    print(" block",b,"ops incl. recompute",tot)
