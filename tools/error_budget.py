"""Where the fp16-operand error of a SegmentSumMLP score comes from: an fp64 simulation of the scoring path on the bench
generator's data (bench.gen_workload) in which each operand rounding of the tcgen05 kernels -- x, W0, the hidden
activation, W1 (row encoder); S, Wl0, y1, Wl1 (head) -- is switched on alone and together.  CPU only:
python tools/error_budget.py.  Output of the committed run: DESIGN.md section 2."""
import numpy as np, torch, sys
sys.path.insert(0, __import__('os').path.dirname(__import__('os').path.dirname(__import__('os').path.abspath(__file__))))
import bench
torch.manual_seed(0)
import vaex_b200
F,H=164,256
net = vaex_b200.SegmentSumMLPModule(F,H,1)
P=4000
feats,sizes,norm,bad = bench.gen_workload(P, 1000, 'cpu')
sd={k:v.double() for k,v in net.state_dict().items()}
X=feats.double(); nv=norm.double()
off=np.concatenate([[0],np.cumsum(sizes)])
seg=torch.from_numpy(np.repeat(np.arange(P),sizes))
def h16(t): return t.float().clamp(-65504,65504).half().double()
def run(rx=False,rw0=False,ra1=False,rw1=False,rs=False,rwl0=False,ry1=False,rwl1=False, split_s=False, split_y1=False):
    W0=sd['segment_encoder.0.weight']/nv[None,:]; b0=sd['segment_encoder.0.bias']
    W1=sd['segment_encoder.2.weight']; b1=sd['segment_encoder.2.bias']
    x = h16(X) if rx else X
    w0 = h16(W0) if rw0 else W0
    a1 = torch.relu(x@w0.T+b0)
    a1 = h16(a1) if ra1 else a1
    w1 = h16(W1) if rw1 else W1
    h2 = torch.relu(a1@w1.T+b1)
    S = torch.zeros(P,H,dtype=torch.double).index_add_(0,seg,h2)
    Wl0=sd['l0.0.weight']; Wl1=sd['l1.0.weight']
    wl0 = h16(Wl0) if rwl0 else Wl0
    wl1 = h16(Wl1) if rwl1 else Wl1
    s_op = h16(S) if rs else S
    if split_s: s_op = h16(S) + h16(S-h16(S))
    y1 = torch.relu(s_op@wl0.T + sd['l0.0.bias']) + S
    y1_op = h16(y1) if ry1 else y1
    if split_y1: y1_op = h16(y1)+h16(y1-h16(y1))
    y2 = torch.relu(y1_op@wl1.T + sd['l1.0.bias']) + y1
    return (y2@sd['decoder.weight'].T + sd['decoder.bias']).squeeze(-1), S
ref,S = run()
ok = np.ones(P,bool); ok[bad]=False
rms = ref[ok].pow(2).mean().sqrt().item()
print('score rms',rms,'S rms',S.pow(2).mean().sqrt().item(), 'max S', S.max().item())
def rep(name, **kw):
    got,_ = run(**kw)
    d=(got-ref)[ok]
    print('%-28s l2 %.2e  worst/rms %.2e  worst rel-elem %.2e'%(name, (d.norm()/ref[ok].norm()).item(), d.abs().max().item()/rms, (d.abs()/ref[ok].abs()).max().item()))
rep('x', rx=True); rep('w0', rw0=True); rep('a1', ra1=True); rep('w1', rw1=True)
rep('S', rs=True); rep('wl0', rwl0=True); rep('y1', ry1=True); rep('wl1', rwl1=True)
rep('segsum all', rx=True,rw0=True,ra1=True,rw1=True)
rep('head all', rs=True,rwl0=True,ry1=True,rwl1=True)
rep('all', rx=True,rw0=True,ra1=True,rw1=True,rs=True,rwl0=True,ry1=True,rwl1=True)
rep('all, split S', rx=True,rw0=True,ra1=True,rw1=True,split_s=True,rwl0=True,ry1=True,rwl1=True)
rep('all, split S,y1', rx=True,rw0=True,ra1=True,rw1=True,split_s=True,rwl0=True,split_y1=True,rwl1=True)
