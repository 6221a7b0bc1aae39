"""check_tcs.py -- one VLUT_ALGO_TCS call per process against the scalar anchor kernel, for a list of env settings."""
import os, subprocess, sys
CHILD = r'''
import sys, torch
sys.path.insert(0, ".")
import vlut_b200
from vlut_b200 import api
K, N, T = int(sys.argv[1]), int(sys.argv[2]), int(sys.argv[3])
ctx = api.Context(0)
g = torch.Generator(device="cuda"); g.manual_seed(1)
gg = torch.randint(0, 81, (api.nbytes(42, K, N),), dtype=torch.uint8, device="cuda", generator=g)
W = ctx.upload(42, K, N, gg)
x = torch.randn((T, K), device="cuda", generator=g)
a = ctx.mul_mat(W, x, algo=api.ALGO_SIMPLE).clone()
torch.cuda.synchronize()
b = ctx.mul_mat(W, x, algo=api.ALGO_TCS).clone()
torch.cuda.synchronize()
print("equal", bool(torch.equal(a, b)))
'''
cases = [("512 128 16", {}), ("512 128 16", {"VLUT_TCS_ISSUERS": "1"}), ("256 128 16", {}), ("4096 128 16", {"VLUT_TCS_ISSUERS": "1"}),
         ("4096 128 16", {}), ("4096 14336 16", {"VLUT_TCS_ISSUERS": "1"}), ("4096 14336 16", {}), ("4096 128 16", {"VLUT_TC_FAST": "0"})]
for args, env in cases:
    e = dict(os.environ); e.update(env); e["CUDA_LAUNCH_BLOCKING"] = "1"
    r = subprocess.run([sys.executable, "-c", CHILD] + args.split(), env=e, capture_output=True, text=True, timeout=120)
    out = (r.stdout.strip().splitlines() or ["-"])[-1]
    err = [l for l in r.stderr.splitlines() if "rror" in l][:1]
    print(args, env, "->", out, err, flush=True)
