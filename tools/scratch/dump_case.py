import os, sys
sys.path.insert(0, "/root/repo")
from asset_asrl_b200 import workloads as W
mode, n, tag = sys.argv[1], int(sys.argv[2]), sys.argv[3]
ph = W.make_phase("lowthrust", mode, n)
nlp = ph.transcribe()
info = nlp.fused_info(); print(info)
for p in range(int(info["pieces"])):
    open(f"/tmp/{tag}_{p}.ptx", "w").write(nlp.fused_ptx(p))
