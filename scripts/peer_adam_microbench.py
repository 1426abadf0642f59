"""One-GPU microbenchmark of the multi-GPU optimiser kernels (mne_grad_push + mne_adam_step_peers) at room0 size: `world` ranks
emulated on one device (every "peer" buffer lives on this GPU), realistic sparsity (half of the map ever touched, in 16 KB
blocks; a third of the touched texels non-zero this step).  Times the kernels of emulated rank 0 only; no NVLink involved, so
what is measured is the kernels' own HBM efficiency."""
import ctypes as C
import os
import sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from mneslam_b200 import _lib, ops

world = int(os.environ.get('WORLD', '2'))
chunk = 1024
n_used = 32339008 + 6208
q = chunk * world
n_total = (n_used + q - 1) // q * q
lib, st = _lib.lib(), ops._stream()
torch.manual_seed(0)
dev = 'cuda'
local = n_total // world
params = [torch.randn(n_total, device=dev) for _ in range(world)]
for p in params[1:]:
    p.copy_(params[0])
grads = [torch.zeros(n_total, device=dev) for _ in range(world)]
dirty = [torch.zeros(world * (local // 32), dtype=torch.uint8, device=dev) for _ in range(world)]
inbox = [torch.zeros(world * local, device=dev) for _ in range(world)]
m0, v0 = torch.zeros(local, device=dev), torch.zeros(local, device=dev)
touched0 = torch.zeros(local // 4, dtype=torch.uint8, device=dev)
pp = (C.c_void_p * world)(*[x.data_ptr() for x in params])
gp = (C.c_void_p * world)(*[x.data_ptr() for x in grads])
dp = (C.c_void_p * world)(*[x.data_ptr() for x in dirty])
ip = (C.c_void_p * world)(*[x.data_ptr() for x in inbox])
seg_ends = (C.c_int64 * 2)(32339008, n_used)
lr, eps, wd = (C.c_float * 2)(5e-3, 1e-2), (C.c_float * 2)(1e-15, 1e-8), (C.c_float * 2)(0.0, 1e-6)

blocks = torch.rand(n_total // 4096, device=dev) < 0.5           # regions of the map that are ever observed


def fill_gradients():
    for r in range(world):
        tex = (torch.rand(n_total // 32, device=dev) < 0.34) & blocks.repeat_interleave(128)
        g = grads[r].view(-1, 32)
        g.zero_()
        g[tex] = torch.randn(int(tex.sum()), 32, device=dev)
        grads[r][n_used:] = 0


def timed(fn):
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize()
    a.record(); fn(); b.record()
    torch.cuda.synchronize()
    return a.elapsed_time(b) * 1e3


t_push, t_adam, t_flag = [], [], []
for t in range(1, 9):
    fill_gradients()
    for r in range(1, world):
        _lib.check(lib.mne_grad_push(grads[r].data_ptr(), n_total, n_total, chunk, world, r, dp, ip, st), 'push')
    if t % 2:      # flags from the gradient buffer itself (a pass over 129 MB) ...
        t_push.append(timed(lambda: _lib.check(lib.mne_grad_push(grads[0].data_ptr(), n_total, n_total, chunk, world, 0, dp, ip, st), 'push')))
    else:          # ... or set by the backward pass (here: derived beforehand), a pass over the 1 MB map
        fl = (grads[0].view(-1, 32) != 0).any(1).to(torch.uint8)
        t_flag.append(timed(lambda: _lib.check(lib.mne_dirty_push_data(fl.data_ptr(), grads[0].data_ptr(), n_total, n_total, chunk, world, 0, dp, ip, st), 'push_data')))
    t_adam.append(timed(lambda: _lib.check(lib.mne_adam_step_peers(world, 0, pp, gp, n_total, chunk, m0.data_ptr(), v0.data_ptr(), 2, seg_ends,
                                                                   lr, eps, wd, 0.9, 0.99, None, t, touched0.data_ptr(), None, None,
                                                                   dirty[0].data_ptr(), inbox[0].data_ptr(), st), 'peers')))
    for r in range(1, world):        # the other ranks consume their share, so that every gradient buffer is clean again
        mr, vr = torch.zeros(local, device=dev), torch.zeros(local, device=dev)
        _lib.check(lib.mne_adam_step_peers(world, r, pp, gp, n_total, chunk, mr.data_ptr(), vr.data_ptr(), 2, seg_ends, lr, eps, wd,
                                           0.9, 0.99, None, t, None, None, None, dirty[r].data_ptr(), inbox[r].data_ptr(), st), 'peers')
        del mr, vr
touched = float(touched0.float().mean())
own4 = local // 4
# bytes of rank 0's optimiser kernel: flags, p / m / v in + m / v out + p out to every rank for the touched groups, the flagged
# gradients (own buffer: read + cleared; inbox: read)
nz_own = float((dirty[0][:local // 32] != 0).float().mean())
adam_bytes = own4 * (1 + world / 8) + own4 * touched * 16 * (5 + world) + own4 * nz_own * 16 * (2 + (world - 1))
print('world %d  grad_push us %s   dirty_push_data us %s' % (world, [round(x, 1) for x in t_push[1:]], [round(x, 1) for x in t_flag[1:]]))
print('world %d  adam_peers us %s' % (world, [round(x, 1) for x in t_adam[2:]]))
print('touched %.3f  flagged %.3f  adam bytes %.1f MB -> %.0f GB/s at the median' % (
    touched, nz_own, adam_bytes / 1e6, adam_bytes / (sorted(t_adam[2:])[len(t_adam[2:]) // 2] * 1e-6) / 1e9))
