"""Per-tensor gradient error of the CUDA path against the reference goldens (max-norm rel and cosine).
Run on a GPU box: python scripts/grad_margins.py"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
from tests.cases import Case
from tests.gpu_util import build_model, cosine, flat_planes, rel, total_loss

for name in ('small_onegrid', 'small_twogrid', 'r0_256', 'sn_128'):
    case = Case(name)
    g, cfg = case.g, case.cfg
    model = build_model(cfg, case.bounding_box, case.planes, case.weights)
    ret = model.forward(case.t('rays_o').cuda(), case.t('rays_d').cuda(), case.t('target_rgb').cuda(),
                        case.t('target_d').cuda(), noise=case.t('noise').cuda())
    total_loss(cfg, ret).backward()
    rows = []
    for w, k in zip(model.decoder.weight_list(), ('sdf0', 'sdf1', 'col0', 'col1')):
        rows.append((k, rel(w.grad.cpu().numpy(), g['g_' + k]), cosine(w.grad.cpu().numpy(), g['g_' + k])))
    for i, p in enumerate(flat_planes(model)):
        got = p.grad.cpu()
        if f'g_plane{i}' in g:
            rows.append((f'plane{i}', rel(got.numpy(), g[f'g_plane{i}']), cosine(got.numpy(), g[f'g_plane{i}'])))
        else:
            vals = got.contiguous().reshape(-1)[torch.from_numpy(g[f'gp_idx{i}'])].numpy()
            if np.abs(g[f'gp_val{i}']).max() > 0:
                rows.append((f'plane{i}(probe)', rel(vals, g[f'gp_val{i}']), cosine(vals, g[f'gp_val{i}'])))
    worst = max(r[1] for r in rows)
    print('%-14s two_grid=%d  worst rel %.2e  min cos %.7f' % (name, not cfg['grid']['oneGrid'], worst, min(r[2] for r in rows)))
    for k, r, c in rows:
        print('    %-16s rel %.2e  cos %.7f' % (k, r, c))
