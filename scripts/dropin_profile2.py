"""Host-side profile of the drop-in iteration with mneslam_b200.optim.Adam (one-line optimiser swap)."""
import cProfile, os, pstats, sys, time
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


def step():
    ro, rd, rgb, d = db.sample(2048)
    ret = model.forward(ro, rd, rgb, d)
    loss = (lw[0] * ret['rgb_loss'] + lw[1] * ret['depth_loss'] + lw[4] * ret['e_fs_loss'] + lw[5] * ret['e_center_loss']
            + lw[6] * ret['e_tail_loss'])
    loss.backward()
    opt.step()
    opt.zero_grad()
    return loss


for _ in range(10):
    step()
torch.cuda.synchronize()
t0 = time.perf_counter()
for _ in range(100):
    step()
torch.cuda.synchronize()
print('ms per step: %.3f' % ((time.perf_counter() - t0) * 1e3 / 100))
a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
torch.cuda._sleep(200000000)          # ~0.1 s of device-side delay: the host runs ahead, the events then bracket pure GPU time
a.record()
for _ in range(50):
    step()
b.record()
torch.cuda.synchronize()
print('GPU ms per step (host ahead): %.3f' % (a.elapsed_time(b) / 50))
pr = cProfile.Profile()
pr.enable()
for _ in range(50):
    step()
pr.disable()
torch.cuda.synchronize()
pstats.Stats(pr).sort_stats('tottime').print_stats(22)
