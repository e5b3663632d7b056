"""Cycle counters of segsum_tc.cu built with -DSEGSUM_PROF: python tools/prof_segsum.py <lib.so>  (bench workload, row image)"""
import ctypes, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from vaex_b200 import _lib
_lib.LIB_PATH = os.path.abspath(sys.argv[1])
import bench, vaex_b200
dev = torch.device('cuda', 0)
P = 1_000_000
feats, sizes, norm, bad = bench.gen_workload(P, 1000, dev)
rows = feats.shape[0]
torch.manual_seed(0)
net = vaex_b200.SegmentSumMLPModule(164, 256, 1).to(dev).eval()
model = net.packed(norm.cpu())
lib = _lib.load(); st = torch.cuda.current_stream().cuda_stream
image = torch.empty(int(lib.vaex_rows_image_bytes(rows)), dtype=torch.uint8, device=dev)
_lib.check(lib.vaex_rows_pack_image(_lib.ptr(feats), rows, 164, _lib.ptr(image), st))
sizes_d = torch.from_numpy(sizes).to(dev); offsets = torch.empty(P + 1, dtype=torch.int32, device=dev)
_lib.check(lib.vaex_segment_offsets(_lib.ptr(sizes_d), P, _lib.ptr(offsets), st))
seg = torch.empty(P, 256, device=dev)
for _ in range(3):
    _lib.check(lib.vaex_segsum_encode(model.h, _lib.ptr(image), _lib.ptr(offsets), P, rows, _lib.ptr(seg), _lib.F_FEATS_IMAGE, st))
torch.cuda.synchronize()
buf = (ctypes.c_ulonglong * 32)()
lib.vaex_debug_segsum_prof(buf)
v = list(buf); n = max(v[0], 1)
print('tiles of cluster 0: %d; issuer loop %.0f clk per tile' % (v[0], v[6] / n))
print('issuer waits per tile: A0_FULL %.0f | ACC1_EMPTY0 %.0f | A1_FULL0 %.0f | A1_FULL1 %.0f | ACC1_EMPTY1 %.0f' % tuple(v[i] / n for i in (1, 2, 3, 4, 5)))
print('epi 0 (warp 8): half 0 wait %.0f work %.0f | half 1 wait %.0f work %.0f' % tuple(v[i] / n for i in (8, 9, 10, 11)))
print('epi 1 pollers: block 0 wait %.0f drain %.0f | block 1 wait %.0f drain %.0f  (image mode: two pollers per block add up)' % tuple(v[i] / n for i in (12, 13, 14, 15)))
