import os, sys
sys.path.insert(0, "/root/repo")
from asset_asrl_b200 import workloads as W
ph = W.make_phase("lowthrust", "LGL7", 90, control_mode="HighestOrderSpline")
nlp = ph.transcribe()
info = nlp.fused_info()
for p in range(int(info["pieces"])):
    open(f"/tmp/hos_{os.environ.get('ASSET_B200_FUSED_DIRECT','1')}_{p}.ptx", "w").write(nlp.fused_ptx(p))
