import sys, os, json, ctypes
sys.path.insert(0, '/root/repo')
import torch
from vaex_b200 import _lib
_lib.LIB_PATH = os.path.abspath('tools/_ab/libvaex_prof.so')
import bench, vaex_b200
dev = torch.device('cuda', 0)
vae, reg, mean, scale, ext, last = bench.extent_models(dev)
lib = _lib.load()
n = 1_000_000
z = torch.randn(n, 64, device=dev); x = torch.randn(n, 48, device=dev)
buf = (ctypes.c_ulonglong * 64)()
def run(name, fn):
    fn(); fn(); torch.cuda.synchronize()
    lib.vaex_debug_chain_prof(buf, 1)
    fn(); torch.cuda.synchronize()
    lib.vaex_debug_chain_prof(buf, 0)
    v = list(buf)
    print(name, 'tiles', v[51], 'cycles/tile', v[50] / max(v[51] - 1, 1))
    for l in range(4):
        for nh in range(2):
            c = v[l * 8 + nh * 4 + 3]
            if c: print('  worker l%d nh%d: wait %.0f  tmem_ld %.0f  math+sts+publish %.0f  (n=%d)' % (l, nh, v[l*8+nh*4]/c, v[l*8+nh*4+1]/c, v[l*8+nh*4+2]/c, c))
        c = max(v[51] - 2, 1)
        if v[l*8+4]: print('  worker l%d: loop top (fetch) -> wait %.0f   loop top -> end of out epilogue %.0f' % (l, v[l*8+4]/v[l*8+3], v[l*8+5]/v[l*8+3]))
        for kh in range(2):
            if v[40 + l * 2 + kh]: print('  issuer l%d kh%d: wait IN %.0f' % (l, kh, v[40 + l*2+kh]/c))
run('decode', lambda: vae.decode(z))
run('predict', lambda: reg.predict_cost(z))
