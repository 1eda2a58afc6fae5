cat > /tmp/one.py <<'PY'
import os, sys
sys.path.insert(0, "/root/repo")
import torch
from rtseg_b200 import modules
from tools.layer_bench import nhwc_rand
m = modules.DSConvBNAct(128, 128, 3, 1).eval().cuda(); x = nhwc_rand((32, 128, 128, 256))
with torch.no_grad():
    for _ in range(3): m(x)
torch.cuda.synchronize()
PY
timeout 600 ncu --set full --clock-control none --import-source on -k regex:dwpw --launch-skip 2 --launch-count 1 -o gpurun_out/dwpw_r02 -f python /tmp/one.py > gpurun_out/ncu31.log 2>&1; echo rc=$?; tail -3 gpurun_out/ncu31.log
