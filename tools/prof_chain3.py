"""Cycle counters of chain_tc3.cu built with -DCHAIN3_PROF (see the kernel): python tools/prof_chain3.py <lib.so>"""
import ctypes, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from vaex_b200 import _lib
_lib.LIB_PATH = os.path.abspath(sys.argv[1])
import bench, vaex_b200
dev = torch.device('cuda', 0)
vae, reg, mean, scale, ext, last = bench.extent_models(dev)
lib = _lib.load()
n = 1_000_000
z = torch.randn(n, 64, device=dev); x = torch.randn(n, 48, device=dev)
buf = (ctypes.c_ulonglong * 64)()
def run(name, fn, layers):
    fn(); fn(); torch.cuda.synchronize()
    lib.vaex_debug_chain3_prof(buf)
    v = list(buf)
    d = max(v[1], 1)
    print('%s: %d tile pairs, %.0f clk per pair' % (name, v[1], v[0] / d))
    print('  issuer waits per pair: DRAIN X %.0f  DRAIN Y %.0f  W_FULL %.0f  IN X %.0f  IN Y %.0f' % (v[2]/d, v[3]/d, v[4]/d, v[5]/d, v[6]/d))
    print('  issuer wait per layer (W_FULL + IN, both tiles):', ['%.0f' % (v[8 + l] / d) for l in range(layers)])
    print('  worker X warp 0: ACC wait per layer', ['%.0f' % (v[16 + l] / d) for l in range(layers)], ' hidden epilogue', ['%.0f' % (v[24 + l] / d) for l in range(layers)])
    print('  worker X warp 0: out layer ld+drain %.0f  pack_input %.0f' % (v[33] / d, v[32] / d))
run('decode', lambda: vae.decode(z), 4)
run('predict', lambda: reg.predict_cost(z), 2)
run('encode', lambda: vae.encode(x), 4)
