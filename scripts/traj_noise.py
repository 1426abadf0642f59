"""How far do five fused iterations drift from autograd + torch.optim.Adam?  (run-to-run spread of the bulk error)"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from mneslam_b200 import synthetic
from mneslam_b200.mapping import FusedMapper
from tests.gpu_util import build_model, flat_planes, total_loss
from tests.test_gpu_mapping import _small_model

for one_grid in (True, False):
    for rep in range(4):
        cfg, bb, planes, weights = _small_model(one_grid=one_grid)
        a = build_model(cfg, bb, planes, weights)
        b = build_model(cfg, bb, planes, weights, as_params=False)
        mp = cfg['mapping']
        groups = [{'params': a.decoder.parameters(), 'weight_decay': 1e-6, 'lr': mp['lr_decoder']},
                  {'params': flat_planes(a)[:6], 'eps': 1e-15, 'lr': mp['lr_embed']}]
        if not one_grid:
            groups.append({'params': flat_planes(a)[6:], 'eps': 1e-15, 'lr': mp['lr_embed_color']})
        opt = torch.optim.Adam(groups, betas=(0.9, 0.99))
        mapper = FusedMapper(b)
        n, s = 200, synthetic.n_samples_total(cfg)
        for it in range(5):
            batch = {k: v.cuda() for k, v in synthetic.make_ray_batch(cfg, n, seed=50 + it).items()}
            noise = synthetic.make_noise(n, s, seed=it).cuda()
            ret = a.forward(batch['rays_o'], batch['rays_d'], batch['target_rgb'], batch['target_d'], noise=noise)
            total_loss(cfg, ret).backward()
            opt.step(); opt.zero_grad()
            mapper.step(batch['rays_o'], batch['rays_d'], batch['target_rgb'], batch['target_d'], noise=noise)
        def err(x, y):
            x, y = x.detach().double().reshape(-1), y.detach().double().reshape(-1)
            return float((x - y).norm() / (y.norm() + 1e-30))
        pe = max(err(pb, pa) for pa, pb in zip(flat_planes(a), flat_planes(b)))
        we = [err(wb, wa) for wa, wb in zip(a.decoder.weight_list(), b.decoder.weight_list())]
        print('one_grid', one_grid, 'rep', rep, 'planes %.2e' % pe, 'weights', ['%.2e' % e for e in we])
