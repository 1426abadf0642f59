"""Two-rank check of the ray-sharded mapping iteration (run under torchrun with 2 GPUs):
the parameters after a few data-parallel iterations equal those of one GPU working on the whole batch."""
import os
import sys

import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

from mneslam_b200 import configs, synthetic                    # noqa: E402
from mneslam_b200.distributed import shard_rays                # noqa: E402
from mneslam_b200.mapping import FusedMapper                   # noqa: E402
from oracle import joint_encoding as J                         # noqa: E402
from tests.gpu_util import build_model, flat_planes            # noqa: E402


def main():
    rank, world, local = int(os.environ['RANK']), int(os.environ['WORLD_SIZE']), int(os.environ['LOCAL_RANK'])
    torch.cuda.set_device(local)
    dist.init_process_group('nccl', device_id=torch.device('cuda', local))
    cfg = configs.small_test_config(n_samples_d=13, n_range_d=7)
    bound = J.enlarged_bound(cfg)
    bb = torch.tensor(cfg['mapping']['bound'], dtype=torch.float64)
    planes, weights = J.init_planes(cfg, bound, seed=2, std=0.3), J.init_weights(cfg, seed=2)
    model = build_model(cfg, bb, planes, weights, as_params=False)
    mapper = FusedMapper(model, world_size=world)
    single = None
    if rank == 0:
        ref_model = build_model(cfg, bb, planes, weights, as_params=False)
        single = FusedMapper(ref_model)
    n, s = 512 + 3, synthetic.n_samples_total(cfg)
    for it in range(4):
        batch = {k: v.cuda() for k, v in synthetic.make_ray_batch(cfg, n, seed=40 + it, zero_depth_frac=0.1).items()}
        noise = synthetic.make_noise(n, s, seed=it).cuda()
        lo, hi = shard_rays(n, rank, world)
        # norm_rays must be the GLOBAL ray count even though the shards differ in size by one ray
        mapper_losses = _step_shard(mapper, batch, noise, lo, hi, n)
        if single is not None:
            ref_losses = single.step(batch['rays_o'], batch['rays_d'], batch['target_rgb'], batch['target_d'], noise=noise)
            assert torch.allclose(mapper_losses[:7], ref_losses[:7], rtol=1e-4, atol=1e-7, equal_nan=True), (mapper_losses, ref_losses)
    torch.cuda.synchronize()
    ok = True
    if single is not None:
        for a, b in zip(flat_planes(model) + model.decoder.weight_list(), flat_planes(ref_model) + ref_model.decoder.weight_list()):
            err = float((a.double() - b.double()).norm() / (b.double().norm() + 1e-30))
            ok = ok and err < 2e-3
            assert err < 2e-3, err
    # every rank holds the same parameters
    flat = mapper.flat_param.clone()
    dist.broadcast(flat, src=0)
    assert torch.equal(flat, mapper.flat_param), 'ranks diverged'
    assert mapper.barrier_timeouts() == 0, 'a cross-rank barrier gave up waiting'
    dist.barrier()
    if rank == 0:
        print('multigpu_check ok')
    dist.destroy_process_group()


def _step_shard(mapper, batch, noise, lo, hi, n_global):
    from mneslam_b200 import ops
    # FusedMapper derives norm_rays from its local batch x world size; with ragged shards pass the global count explicitly
    mapper.global_rays_override = n_global
    return mapper.step(batch['rays_o'][lo:hi].contiguous(), batch['rays_d'][lo:hi].contiguous(),
                       batch['target_rgb'][lo:hi].contiguous(), batch['target_d'][lo:hi].contiguous(),
                       noise=noise[lo:hi].contiguous())


if __name__ == '__main__':
    main()
