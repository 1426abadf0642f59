"""Cost of the dirty-texel marking inside the backward kernel (one GPU; events around mne_render_backward)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from mneslam_b200 import JointEncoding, configs, ops, synthetic
from mneslam_b200.mapping import DeviceKeyframeRays, FusedMapper

cfg = configs.replica_room0()
torch.manual_seed(0)
model = JointEncoding(cfg, torch.tensor(cfg['mapping']['bound'], dtype=torch.float64)).cuda()
mapper = FusedMapper(model)
rays, poses = synthetic.make_keyframe_db(cfg, n_kf=20, seed=0)
db = DeviceKeyframeRays(rays.cuda(), poses.cuda())
for _ in range(20):
    mapper.step(*db.sample(2048))
dirty = torch.zeros((mapper.flat_grad.numel() + 31) // 32, dtype=torch.uint8, device='cuda')


def run(mark, iters=30):
    mapper.dirty_local = dirty if mark else None
    ts = []
    for _ in range(iters):
        ro, rd, rgb, d = db.sample(2048)
        cfgr = ops.make_render_cfg(cfg, norm_rays=2048)
        tables = ops.SampleTables.get(cfg, 'cuda', True)
        s = tables[0].numel() + tables[1].numel()
        acc = ops.new_acc('cuda')
        z = ops.sample_rays(ops.make_sample_cfg(tables, True), cfgr, 2048, s, d.reshape(-1), torch.rand(2048, s, device='cuda'), acc, 'cuda')
        out = ops.render_forward(mapper._scene(False), cfgr, ro, rd, z, rgb, d.reshape(-1), acc, want_aux=False)
        ops.loss_finalize(cfgr, s, acc)
        scr = ops.backward_scratch(2048, s, 'cuda')
        sc = mapper._scene(True)
        torch.cuda.synchronize()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        ops.render_backward(sc, cfgr, ro, rd, z, out['raw'], rgb, d.reshape(-1), acc, loss_grads=mapper.loss_grads, scratch=scr)
        b.record(); torch.cuda.synchronize()
        ts.append(a.elapsed_time(b) * 1e3)
        mapper.flat_grad.zero_(); dirty.zero_()
    ts.sort()
    return round(ts[len(ts) // 2], 1)


print('backward us  no marking:', run(False), ' marking:', run(True), ' no marking:', run(False), ' marking:', run(True))
