"""Event-timed breakdown of the multi-GPU optimiser step (run under torchrun)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch, torch.distributed as dist
from mneslam_b200 import JointEncoding, _lib, configs, ops, synthetic
from mneslam_b200.mapping import DeviceKeyframeRays, FusedMapper

rank, world, local = int(os.environ['RANK']), int(os.environ['WORLD_SIZE']), int(os.environ['LOCAL_RANK'])
torch.cuda.set_device(local)
os.environ.setdefault('NCCL_DEBUG', 'WARN')
dist.init_process_group('nccl', device_id=torch.device('cuda', local))
cfg = configs.replica_room0()
torch.manual_seed(0)
model = JointEncoding(cfg, torch.tensor(cfg['mapping']['bound'], dtype=torch.float64)).cuda()
mapper = FusedMapper(model, world_size=world)
rays, poses = synthetic.make_keyframe_db(cfg, n_kf=20, seed=0)
db = DeviceKeyframeRays(rays.cuda(), poses.cuda())
for _ in range(30):
    mapper.step(*db.sample(2048))
torch.cuda.synchronize(); dist.barrier()
lib, st, pr = _lib.lib(), ops._stream(), mapper.peers
n_seg, ends, lr, eps, wd = mapper._seg_head
def hand_over():
    if mapper.dirty_local is not None and mapper.push_gradients:
        lib.mne_dirty_push_data(mapper.dirty_local.data_ptr(), mapper.flat_grad.data_ptr(), mapper.flat_grad.numel(), mapper.dense_from, 1024,
                                world, rank, pr['dirty_ptrs'], pr['inbox_ptrs'], st)
    elif mapper.dirty_local is not None:
        lib.mne_dirty_push(mapper.dirty_local.data_ptr(), mapper.flat_grad.numel(), mapper.dense_from, 1024, world, rank, pr['dirty_ptrs'], st)
    elif mapper.push_gradients:
        lib.mne_grad_push(mapper.flat_grad.data_ptr(), mapper.flat_grad.numel(), mapper.dense_from, 1024, world, rank, pr['dirty_ptrs'], pr['inbox_ptrs'], st)
    else:
        lib.mne_grad_dirty_map(mapper.flat_grad.data_ptr(), mapper.flat_grad.numel(), 1024, world, rank, pr['dirty_ptrs'], st)
names = ['fwd_bwd', 'dirty', 'barrier0', 'adam_peers', 'barrier1']
tot = {k: 0.0 for k in names}
K = 20
for it in range(K):
    ev = [torch.cuda.Event(enable_timing=True) for _ in range(6)]
    batch = db.sample(2048)
    ev[0].record()
    mapper.forward_backward(*batch)
    ev[1].record()
    hand_over()
    ev[2].record()
    pr['hg'].barrier(channel=0)
    ev[3].record()
    lib.mne_adam_tick(mapper.step_dev.data_ptr(), st)
    lib.mne_adam_step_peers(world, mapper.shards.rank, pr['param_ptrs'], pr['grad_ptrs'], mapper.flat_grad.numel(), 1024, mapper.flat_m.data_ptr(),
                            mapper.flat_v.data_ptr(), n_seg, ends, lr, eps, wd, 0.9, 0.99, mapper.step_dev.data_ptr(), 0,
                            ops._ptr(mapper.touched), pr['mc_param'] or None, None, pr['dirty'].data_ptr(), pr['inbox'].data_ptr() if mapper.push_gradients else None, st)
    ev[4].record()
    pr['hp'].barrier(channel=1)
    ev[5].record()
    torch.cuda.synchronize()
    for k, a, b in zip(names, ev[:-1], ev[1:]):
        tot[k] += a.elapsed_time(b) * 1e3 / K
dirty_frac = pr['dirty'].float().mean().item()
print('rank %d  us: %s  dirty texels this step: %.3f  touched groups: %.3f' % (rank, {k: round(v, 1) for k, v in tot.items()}, dirty_frac,
      mapper.touched.float().mean().item()), flush=True)
# where does the peer kernel's time go?  Re-run it with some of the peer pointers replaced by local ones (the values
# are garbage afterwards -- this is the last thing the script does)
import ctypes as C
def ptrs(kind, local_only):
    own = [int(x) for x in pr[kind]]
    return (C.c_void_p * world)(*([own[rank]] * world if local_only else own))
def timed(param_local, grad_local):
    pp, gp, dp = ptrs('param_ptrs', param_local), ptrs('grad_ptrs', grad_local), pr['dirty'].data_ptr()
    ts = []
    for it in range(6):
        mapper.forward_backward(*db.sample(2048))
        hand_over()
        torch.cuda.synchronize(); dist.barrier()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        lib.mne_adam_step_peers(world, rank, pp, gp, mapper.flat_grad.numel(), 1024, mapper.flat_m.data_ptr(),
                                mapper.flat_v.data_ptr(), n_seg, ends, lr, eps, wd, 0.9, 0.99, mapper.step_dev.data_ptr(), 0,
                                ops._ptr(mapper.touched), None, None, dp, pr['inbox'].data_ptr() if mapper.push_gradients else None, st)
        b.record(); torch.cuda.synchronize(); dist.barrier()
        ts.append(a.elapsed_time(b) * 1e3)
    return round(sorted(ts)[len(ts) // 2], 1)
res = {'all_peer': timed(False, False), 'param_local': timed(True, False), 'all_local': timed(True, True)}
print('rank %d  adam_peers variants us: %s' % (rank, res), flush=True)
dist.barrier()
os._exit(0)
