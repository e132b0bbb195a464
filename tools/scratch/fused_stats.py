import os, sys, time
sys.path.insert(0, "/root/repo")
from asset_asrl_b200 import workloads as W
from asset_asrl_b200.Solvers import NonLinearProgram as NLP
mode = sys.argv[1] if len(sys.argv) > 1 else "LGL5"
ph = W.make_phase("lowthrust", mode, 256)
nlp = ph.transcribe()
info = nlp.fused_info()
print(info)
for p in range(int(info["pieces"])):
    ptx = nlp.fused_ptx(p)
    open(f"/tmp/fusedD_{mode}_{p}.ptx", "w").write(ptx)
    t=time.time()
    r = NLP.assemble_ptx(ptx)
    print(p, "ptx", len(ptx), "cubin/regs/spill", r, "%.1fs"%(time.time()-t), "v4", ptx.count("st.global.v4"), "v2", ptx.count("st.global.v2"), "s1", ptx.count("st.global.f64"), "parks", ptx.count("st.shared.f64"), "reloads", ptx.count("ld.shared.f64"))
