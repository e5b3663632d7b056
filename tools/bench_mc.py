import sys, json, torch
sys.path.insert(0, '/root/repo')
import bench, vaex_b200
from vaex_b200 import _lib
dev = torch.device('cuda', 0)
vae, reg, *_ = bench.extent_models(dev)
lib = _lib.load(); st = torch.cuda.current_stream().cuda_stream
n = 1_000_000
z = torch.randn(n, 64, device=dev)
mean, var = torch.empty(n, device=dev), torch.empty(n, device=dev)
def timed(fn, it=5):
    fn(); fn()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize(); a.record()
    for _ in range(it): fn()
    b.record(); torch.cuda.synchronize()
    return a.elapsed_time(b) / it
out = {}
for T in (1, 5, 20):
    out['mc_T%d_ms' % T] = timed(lambda: _lib.check(lib.vaex_mc_predict_cost(reg.packed().h, _lib.ptr(z), n, T, None, 777, 0.1, _lib.ptr(mean), _lib.ptr(var), st)))
masks = (torch.rand(5, n, 256, device=dev) >= 0.1).to(torch.uint8)
out['mc_T5_recorded_masks_ms'] = timed(lambda: _lib.check(lib.vaex_mc_predict_cost(reg.packed().h, _lib.ptr(z), n, 5, _lib.ptr(masks), 0, 0.1, _lib.ptr(mean), _lib.ptr(var), st)))
print(json.dumps(out))
