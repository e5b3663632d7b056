"""One decode+snap and one predict_cost launch of the tcgen05 chain on 1M latents (ncu target)."""
import numpy as np
import torch

import vaex_b200

torch.manual_seed(0)
n = 1 << 20
vae = vaex_b200.VAE_feature_head(48, None, 64, 256).cuda().eval()
reg = vaex_b200.VAECostPredictor(48, hidden_dim=256, latent_dim=64).cuda().eval()
z = torch.randn(n, 64, device="cuda")
md = torch.from_numpy(np.random.default_rng(0).uniform(0.5, 3, 48).astype(np.float32)).cuda()
sd = torch.from_numpy(np.random.default_rng(1).uniform(0.3, 1.5, 48).astype(np.float32)).cuda()
for _ in range(3):
    vae.decode(z)
    vae.decode_and_snap(z, md, sd, want_x=False)
    reg.predict_cost(z)
torch.cuda.synchronize()
print("ok")
