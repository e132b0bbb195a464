import sys, os; sys.path.insert(0, os.path.dirname(os.path.abspath(__file__))); sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import asset_asrl_b200 as ast, numpy as np
from ref_assembly import assemble
vf, oc = ast.VectorFunctions, ast.OptimalControl
def cmp(name, a, b):
    sc = max(np.abs(b).max() if b.size else 0, 1e-300)
    e = np.abs(a-b).max() if b.size else 0
    print(f"  {name}: maxabs {e:.3e} rel-to-max {e/sc:.3e}")
    return e/sc
rng = np.random.default_rng(1)
args = oc.ODEArguments(6,3)
r = args.head3(); v = args.segment3(3); u = args.tail3()
O = oc.ODEBase(vf.stack([v, r.normalized_power3()*(-1.0) + u*0.02]),6,3)
worst=0
for mode,CS in [("LGL3",2),("LGL5",3),("LGL7",4),("Trapezoidal",2)]:
    N=7
    S=(CS-1)*N+1
    T=np.zeros((S,10)); th=np.linspace(0,2,S); T[:,0]=np.cos(th); T[:,1]=np.sin(th); T[:,3]=-np.sin(th); T[:,4]=np.cos(th); T[:,6]=th; T[:,7:]=0.5
    T[:,:6]+=0.01*rng.uniform(-1,1,(S,6)); T[:,7:]+=0.1*rng.uniform(-1,1,(S,3))
    ph=O.phase(mode,T,N)
    ph.addBoundaryValue("Front",range(7),T[0,:7])
    ph.addLUNormBound("Path",[7,8,9],0.001,1.0)
    ph.addBoundaryValue("Back",range(6),T[-1,:6])
    ph.addIntegralObjective(vf.Arguments(3).squared_norm()*0.5,[7,8,9])
    ph.addDeltaTimeObjective(1.0)
    nlp=ph.transcribe()
    d=nlp.dims(); print(mode,d)
    X=ph.makeSolverInput()
    LE=100*rng.uniform(-1,1,d["neq"]); LI=100*rng.uniform(-1,1,d["niq"])
    sc=rng.uniform(0.5,1.5,d["n_solver"])
    nlp.to_device(0); nlp.set_solver_coeffs(sc)
    out=nlp.evalKKT(0.7,X,LE,LI)
    ref=assemble(nlp,X,LE,LI,0.7,solver_coeffs=sc)
    for k in ["KKT","FXE","FXI","PGX","AGX"]: worst=max(worst,cmp(k,out[k],ref[k]))
    print("  val",out["val"],ref["val"])
    worst=max(worst,abs(out["val"]-ref["val"])/max(abs(ref["val"]),1e-300))
    print("  launches",nlp.launches())
print("WORST",worst)
assert worst<1e-12
