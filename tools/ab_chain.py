"""A/B of the chain kernels: python tools/ab_chain.py <lib.so | -> (VAEX_CHAIN_V2=1 selects chain_tc.cu) -- prints ms per 1e6 rows (decode, decode+snap, predict, encode)."""
import sys, os, json
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from vaex_b200 import _lib
if len(sys.argv) > 1 and sys.argv[1] != '-':
    _lib.LIB_PATH = os.path.abspath(sys.argv[1])
import bench, vaex_b200
dev = torch.device('cuda', 0)
vae, reg, mean, scale, ext, last = bench.extent_models(dev)
n = 1_000_000
z = torch.randn(n, 64, device=dev); x = torch.randn(n, 48, device=dev)
mean_d, scale_d = torch.from_numpy(mean).to(dev), torch.from_numpy(scale).to(dev)
def timed(fn, it=20):
    for _ in range(3): fn()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize(); a.record()
    for _ in range(it): fn()
    b.record(); torch.cuda.synchronize()
    return round(a.elapsed_time(b) / it, 4)
out = {}
for rep in range(2):
    out['decode%d' % rep] = timed(lambda: vae.decode(z))
    out['decode_snap%d' % rep] = timed(lambda: vae.decode_snap(z, mean_d, scale_d, log1p=True))
    out['predict%d' % rep] = timed(lambda: reg.predict_cost(z))
    out['encode%d' % rep] = timed(lambda: vae.encode(x))
print(os.path.basename(_lib.LIB_PATH), json.dumps(out))
