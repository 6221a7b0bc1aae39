"""profile_quant.py K T -- a few launches of the activation quantise kernel for ncu:
    ncu --set full --import-source on -k regex:quantize -s 2 -c 1 -o out python scripts/profile_quant.py 14336 2048"""
import sys

sys.path.insert(0, ".")
import torch

import vlut_b200  # noqa: F401
from vlut_b200 import api

K, T = int(sys.argv[1]), int(sys.argv[2])
dev = torch.device("cuda", 0)
ctx = api.Context(dev)
type_ = api.TYPE_BY_NAME["I2_V_4" if K % 512 == 0 or K % 16 == 0 else "I1_V_2"]
xs = [torch.randn((T, K), device=dev) for _ in range(4)]
for i in range(4):
    ctx.quantize_act(type_, xs[i])
torch.cuda.synchronize()
ctx.close()
