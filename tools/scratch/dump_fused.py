import os, sys, time
os.environ["ASSET_B200_FUSED_PIECES"] = sys.argv[2] if len(sys.argv) > 2 else "1"
os.environ["ASSET_B200_FUSED_REGS"] = "100000"
sys.path.insert(0, "/root/repo")
from asset_asrl_b200 import workloads as W
mode = sys.argv[1] if len(sys.argv) > 1 else "LGL5"
ph = W.make_phase("lowthrust", mode, 256)
nlp = ph.transcribe()
t=time.time()
info = nlp.fused_info()
print(info, time.time()-t, file=sys.stderr)
np_ = info["pieces"] if isinstance(info, dict) and "pieces" in info else 1
for p in range(int(np_)):
    ptx = nlp.fused_ptx(p)
    open(f"/tmp/fused_{mode}_{p}.ptx", "w").write(ptx)
    print(p, len(ptx), file=sys.stderr)
