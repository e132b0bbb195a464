import sys, collections
sys.path.insert(0,'/root/repo/tools/scratch')
sys.setrecursionlimit(100000)
from sched_sim import parse
srcs,kinds,outs=parse(sys.argv[1])
n=len(srcs)
depth=[0]*n
cone=[0]*n
for i in range(n):
    if kinds[i] in("ld","const"): depth[i]=0
    else: depth[i]=1+max([depth[s] for s in srcs[i]] or [0])
users=[[] for _ in range(n)]
for i in range(n):
    for s in srcs[i]: users[s].append(i)
nuse=[len(users[i]) for i in range(n)]
for o in outs: nuse[o]+=1
print("depth hist", sorted(collections.Counter(depth).items()))
def run(maxdepth):
    # nodes with depth<=maxdepth are rematerialised at every use (never kept live); loads free
    remat=[kinds[i] in("ld","const") or depth[i]<=maxdepth for i in range(n)]
    conecost=[0]*n
    for i in range(n):
        if kinds[i] in("ld","const"): conecost[i]=0
        elif remat[i]: conecost[i]=1+sum(conecost[s] for s in srcs[i])
    emitted=[False]*n; sched=[]; extra=0
    def visit(i):
        nonlocal extra
        if remat[i]:
            return
        if emitted[i]: return
        for s in srcs[i]: visit(s)
        emitted[i]=True; sched.append(i)
    marks=[]
    for o in outs:
        visit(o); marks.append(len(sched))
    total_ops=0
    for i in sched: total_ops+=1+sum(conecost[s] for s in srcs[i])
    for o in outs:
        if remat[o]: total_ops+=conecost[o]
    pos={v:k for k,v in enumerate(sched)}
    lastuse={}
    # uses through remat chains: a kept value used by a remat node is used whenever that remat node is used
    def kept_srcs(i, acc):
        for s in srcs[i]:
            if kinds[s] in("ld","const"): continue
            if remat[s]: kept_srcs(s, acc)
            else: acc.add(s)
    for k,i in enumerate(sched):
        acc=set(); kept_srcs(i,acc)
        for s in acc: lastuse[s]=k
    for oi,o in enumerate(outs):
        if remat[o]:
            acc=set(); kept_srcs(o,acc)
            for s in acc: lastuse[s]=max(lastuse.get(s,-1), marks[oi]-1)
        else: lastuse[o]=max(lastuse.get(o,-1), marks[oi]-1)
    delta=[0]*(len(sched)+2)
    for i in sched:
        d=pos[i]; l=lastuse.get(i,d); delta[d]+=1; delta[l+1]-=1
    c=0;mx=0
    for k in range(len(sched)+1):
        c+=delta[k]; mx=max(mx,c)
    print("remat depth<=",maxdepth,": kept ops",len(sched),"total ops executed",total_ops,"max live",mx)
for d in range(0,8): run(d)
