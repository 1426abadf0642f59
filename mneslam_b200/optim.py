"""``torch.optim``-shaped Adam over the library's fused kernel, for the reference's own optimiser construction.

``create_optimizer`` (mneslam_mp.py:431-469) builds ``torch.optim.Adam`` over the decoder parameters and the plane
``nn.Parameter`` s; swapping the class there (one line) keeps every call site -- ``map_optimizer.step()``,
``zero_grad()``, ``state_dict()`` -- and replaces torch's multi-tensor path (nine launches over the 129 MB map plus
their host-side bookkeeping) by one ``mne_adam_step`` launch per tensor.  Same update as ``torch.optim.Adam`` without
amsgrad / maximize (tested to a few ulp)."""
import torch

from . import _lib, ops


class Adam(torch.optim.Optimizer):
    def __init__(self, params, lr=1e-3, betas=(0.9, 0.999), eps=1e-8, weight_decay=0.0):
        if lr < 0 or eps < 0 or not (0 <= betas[0] < 1 and 0 <= betas[1] < 1) or weight_decay < 0:
            raise ValueError('invalid Adam hyper-parameters')
        super().__init__(params, dict(lr=lr, betas=betas, eps=eps, weight_decay=weight_decay))

    @staticmethod
    def _dense(t):
        return t.is_contiguous() or (t.dim() == 4 and t.is_contiguous(memory_format=torch.channels_last))

    @torch.no_grad()
    def step(self, closure=None):
        loss = None
        if closure is not None:
            with torch.enable_grad():
                loss = closure()
        lib, st = _lib.lib(), ops._stream()
        for group in self.param_groups:
            b1, b2 = group['betas']
            for p in group['params']:
                g = p.grad
                if g is None:
                    continue
                if not (p.is_cuda and p.dtype == torch.float32 and g.dtype == torch.float32):
                    raise TypeError('mneslam_b200.optim.Adam handles float32 CUDA parameters')
                if not self._dense(p):
                    raise TypeError('parameters must be dense (contiguous or channels-last)')
                if g.stride() != p.stride():
                    g = torch.empty_like(p).copy_(g)              # same element order as the parameter
                state = self.state[p]
                if not state:
                    state['step'] = 0
                    state['exp_avg'] = torch.zeros_like(p)        # preserve_format: same strides as p
                    state['exp_avg_sq'] = torch.zeros_like(p)
                state['step'] += 1
                _lib.check(lib.mne_adam_step(p.data_ptr(), g.data_ptr(), state['exp_avg'].data_ptr(),
                                             state['exp_avg_sq'].data_ptr(), p.numel(), group['lr'], b1, b2, group['eps'],
                                             group['weight_decay'], None, state['step'], 0, st), 'mne_adam_step')
        return loss
