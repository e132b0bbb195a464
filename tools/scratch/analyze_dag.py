import re, sys, collections
src = open(sys.argv[1]).read()
body = src[src.index(".entry"):]
ops = []   # (dst, [srcs], kind)
outs = []  # value reg per output in order
prod = {}
reg = re.compile(r"%fd\d+")
for line in body.split("\n"):
    line=line.strip()
    if line.startswith("fma.rn.f64") or line.startswith("div.rn") or line.startswith("sqrt.rn") or line.startswith("rcp.rn") or line.startswith("neg.f64") or line.startswith("mul.rn.f64") or line.startswith("abs.f64"):
        r = reg.findall(line)
        ops.append((r[0], r[1:], line.split()[0])); prod[r[0]] = len(ops)-1
    elif line.startswith("ld.global.nc.f64") or line.startswith("mov.f64"):
        r = reg.findall(line)
        if r:
            ops.append((r[0], r[1:], "ld")); prod[r[0]] = len(ops)-1
    elif line.startswith("st.shared.f64 [%rso"):
        r = reg.findall(line); outs.append(r[0])
    elif line.startswith("{ .param"):
        r = reg.findall(line); ops.append((r[-1], r[:-1], "call")); prod[r[-1]] = len(ops)-1
n = len(ops)
print("ops", n, "outs", len(outs), "kinds", collections.Counter(k for _,_,k in ops))
# first/last output index using each op (transitively)
first = [10**9]*n; last=[-1]*n
users = [[] for _ in range(n)]
for i,(d,s,k) in enumerate(ops):
    for x in s:
        if x in prod: users[prod[x]].append(i)
outuse = collections.defaultdict(list)
for oi,v in enumerate(outs):
    if v in prod: outuse[prod[v]].append(oi)
for i in range(n-1,-1,-1):
    f,l = first[i], last[i]
    for oi in outuse.get(i,[]): f=min(f,oi); l=max(l,oi)
    for u in users[i]: f=min(f,first[u]); l=max(l,last[u])
    first[i],last[i]=f,l
span = [last[i]-first[i] for i in range(n)]
# live across output boundary: number of values with first<=o<last... (value must be alive from its first needed output to last)
live = [0]*(len(outs)+1)
for i in range(n):
    if ops[i][2]=="ld": continue
    for o in range(first[i], last[i]): live[o+1]+=1
print("max values needed across output boundary (excl loads):", max(live), "avg", sum(live)/len(live))
# histogram of span
h = collections.Counter()
for i in range(n):
    if ops[i][2]=="ld": continue
    s=span[i]
    h[0 if s==0 else 1 if s<8 else 2 if s<32 else 3 if s<128 else 4 if s<400 else 5]+=1
print("span hist (0, <8, <32, <128, <400, >=400):", [h[k] for k in range(6)])
# 'core' = values with span>=128: count and how many ops are needed to compute them (closure)
for thr in (32,128,400):
    core=[i for i in range(n) if span[i]>=thr and ops[i][2]!="ld"]
    print("thr",thr,"core values",len(core))
# depth
depth=[0]*n
for i,(d,s,k) in enumerate(ops):
    depth[i]=1+max([depth[prod[x]] for x in s if x in prod] or [0])
print("critical path", max(depth))
# live profile print every 50 outputs
print([live[o] for o in range(0,len(outs),40)])
# fanout distribution
fo=collections.Counter(min(len(users[i])+len(outuse.get(i,[])),20) for i in range(n))
print("fanout", sorted(fo.items()))
