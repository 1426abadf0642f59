"""Host wall time of each call of the drop-in iteration (reference call sites, mneslam_b200.optim.Adam): where the host-bound
0.7 ms goes.  The GPU runs ahead of nothing here (0.33 ms of device work per iteration), so perf_counter deltas are host time."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from mneslam_b200 import JointEncoding, configs, synthetic
from mneslam_b200.mapping import DeviceKeyframeRays, loss_weight_vector
from mneslam_b200.optim import Adam

cfg = configs.replica_room0()
torch.manual_seed(0)
model = JointEncoding(cfg, torch.tensor(cfg['mapping']['bound'], dtype=torch.float64)).cuda()
model.train()
for grp in model.all_planes:
    for i, p in enumerate(grp):
        grp[i] = torch.nn.Parameter(p)
mp = cfg['mapping']
opt = Adam([{'params': model.decoder.parameters(), 'weight_decay': 1e-6, 'lr': mp['lr_decoder']},
            {'params': model._flat_planes(), 'eps': 1e-15, 'lr': mp['lr_embed']}], betas=(0.9, 0.99))
lw = loss_weight_vector(cfg)
rays, poses = synthetic.make_keyframe_db(cfg, n_kf=20, seed=0)
db = DeviceKeyframeRays(rays.cuda(), poses.cuda())
T = {k: 0.0 for k in ('draw', 'forward', 'loss', 'backward', 'step', 'zero_grad')}
pc = time.perf_counter


def step(acc):
    t0 = pc()
    ro, rd, rgb, d = db.sample(2048)
    t1 = pc()
    ret = model.forward(ro, rd, rgb, d)
    t2 = pc()
    loss = (lw[0] * ret['rgb_loss'] + lw[1] * ret['depth_loss'] + lw[4] * ret['e_fs_loss'] + lw[5] * ret['e_center_loss']
            + lw[6] * ret['e_tail_loss'])
    t3 = pc()
    loss.backward()
    t4 = pc()
    opt.step()
    t5 = pc()
    opt.zero_grad()
    t6 = pc()
    if acc:
        for k, v in zip(T, (t1 - t0, t2 - t1, t3 - t2, t4 - t3, t5 - t4, t6 - t5)):
            T[k] += v


for _ in range(20):
    step(False)
torch.cuda.synchronize()
n = 300
t0 = pc()
for _ in range(n):
    step(True)
torch.cuda.synchronize()
print('ms per step: %.3f' % ((pc() - t0) * 1e3 / n))
for k, v in T.items():
    print('%-10s %.1f us' % (k, v * 1e6 / n))
