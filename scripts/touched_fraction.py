"""How much of the map does the bench workload ever touch?  (fraction of texels with a non-zero Adam second moment)"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from mneslam_b200 import JointEncoding, configs, synthetic
from mneslam_b200.mapping import DeviceKeyframeRays, FusedMapper

cfg = configs.replica_room0()
torch.manual_seed(0)
model = JointEncoding(cfg, torch.tensor(cfg['mapping']['bound'], dtype=torch.float64)).cuda()
mapper = FusedMapper(model)
rays, poses = synthetic.make_keyframe_db(cfg, n_kf=20, seed=0)
db = DeviceKeyframeRays(rays.cuda(), poses.cuda())
for it in range(1, 201):
    mapper.step(*db.sample(2048))
    if it in (1, 10, 50, 200):
        v = mapper.flat_v[:mapper.total - 6208].reshape(-1, 32)
        g_touched = (v != 0).any(1).float().mean().item()
        print(it, 'texels ever touched: %.3f' % g_touched, flush=True)
