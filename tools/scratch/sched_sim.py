import re, sys, collections
sys.setrecursionlimit(100000)
def parse(path):
    src = open(path).read()
    body = src[src.index(".entry"):]
    ops = []; outs = []; prod = {}
    reg = re.compile(r"%fd\d+")
    for line in body.split("\n"):
        line=line.strip()
        if line.split(" ")[0] in ("fma.rn.f64","div.rn.f64","sqrt.rn.f64","rcp.rn.f64","neg.f64","mul.rn.f64","abs.f64"):
            r = reg.findall(line); ops.append((r[0], r[1:], line.split()[0])); prod[r[0]] = len(ops)-1
        elif line.startswith("ld.global.nc.f64"):
            r = reg.findall(line); ops.append((r[0], [], "ld")); prod[r[0]] = len(ops)-1
        elif line.startswith("mov.f64"):
            r = reg.findall(line)
            if len(r)==1: ops.append((r[0], [], "const")); prod[r[0]] = len(ops)-1
            else: ops.append((r[0], r[1:], "mov")); prod[r[0]] = len(ops)-1
        elif line.startswith("st.shared.f64 [%rso"):
            r = reg.findall(line); outs.append(prod[r[0]])
        elif line.startswith("{ .param"):
            r = reg.findall(line); ops.append((r[-1], r[:-1], "call")); prod[r[-1]] = len(ops)-1
    srcs = [[prod[x] for x in s if x in prod] for _,s,_ in ops]
    kinds = [k for _,_,k in ops]
    return srcs, kinds, outs
def simulate(srcs, kinds, order_outs, free_loads=True):
    """DFS post-order schedule; returns (max live, live profile at each output, total ops)"""
    n=len(srcs); emitted=[False]*n; sched=[]
    def visit(i):
        if emitted[i]: return
        for s in srcs[i]: visit(s)
        emitted[i]=True; sched.append(i)
    marks=[]
    for o in order_outs:
        visit(o); marks.append(len(sched))
    pos={v:k for k,v in enumerate(sched)}
    lastuse={}
    for k,i in enumerate(sched):
        for s in srcs[i]: lastuse[s]=k
    outpos=collections.defaultdict(int)
    for oi,o in enumerate(order_outs): lastuse[o]=max(lastuse.get(o,-1), marks[oi]-1)
    delta=[0]*(len(sched)+2)
    for i in sched:
        if kinds[i] in ("const",) : continue
        if free_loads and kinds[i]=="ld": continue
        d=pos[i]; l=lastuse.get(i,d)
        delta[d]+=1; delta[l+1]-=1
    live=[];c=0
    for k in range(len(sched)+1):
        c+=delta[k]; live.append(c)
    prof=[live[m-1] for m in marks]
    return max(live), prof, len(sched)
if __name__=="__main__":
    srcs,kinds,outs=parse(sys.argv[1])
    mx,prof,nops=simulate(srcs,kinds,outs)
    print("orig order: max live",mx,"ops",nops, prof[::40])
    mx,prof,nops=simulate(srcs,kinds,outs,free_loads=False)
    print("orig order, loads held: max live",mx)
