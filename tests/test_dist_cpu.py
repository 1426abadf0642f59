"""World-size-2 tests of the multi-GPU host logic on the CPU (gloo): ray sharding, the loss accumulators'
additivity, and the reduce-scatter -> sliced Adam -> all-gather optimiser step.  The CUDA Adam kernel is replaced
by a torch stand-in with the same formulas; everything else is the code the NCCL path runs."""
import os
import socket

import numpy as np
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from mneslam_b200 import configs, synthetic
from mneslam_b200.distributed import FlatShards, shard_rays, sharded_optimizer_step
from oracle import joint_encoding as J


def _free_port():
    s = socket.socket()
    s.bind(('127.0.0.1', 0))
    port = s.getsockname()[1]
    s.close()
    return port


def _torch_adam(param, grad, m, v, step, ends, hparams, betas=(0.9, 0.99)):
    begin = 0
    for end, (lr, eps, wd) in zip(ends, hparams):
        p, g = param[begin:end], grad[begin:end]
        g = g + wd * p
        m[begin:end] += (1 - betas[0]) * (g - m[begin:end])
        v[begin:end] = betas[1] * v[begin:end] + (1 - betas[1]) * g * g
        bc1, bc2 = 1 - betas[0] ** step, 1 - betas[1] ** step
        p -= (lr / bc1) * m[begin:end] / (v[begin:end].sqrt() / bc2 ** 0.5 + eps)
        begin = end


def _worker(rank, world, port, out):
    os.environ.update({'MASTER_ADDR': '127.0.0.1', 'MASTER_PORT': str(port)})
    dist.init_process_group('gloo', rank=rank, world_size=world)
    try:
        torch.manual_seed(0)
        total, seg_ends = 1004, [640, 1004]      # pads to 1008 for two ranks
        hparams = [(5e-3, 1e-15, 0.0), (1e-2, 1e-8, 1e-6)]
        shards = FlatShards(total, world, rank, seg_ends, hparams)
        param0 = torch.randn(shards.padded)
        param0[total:] = 0
        flat_param = param0.clone()
        m = torch.zeros(shards.shard_len)
        v = torch.zeros(shards.shard_len)
        ref_p, ref_m, ref_v = param0[:total].clone(), torch.zeros(total), torch.zeros(total)
        for step in range(1, 4):
            grads = [torch.randn(shards.padded, generator=torch.Generator().manual_seed(100 * step + r)) for r in range(world)]
            for g in grads:
                g[total:] = 0
            flat_grad = grads[rank].clone()
            grad_shard = torch.empty(shards.shard_len)

            def adam_fn(p_slice, g_slice, ends, hp):
                _torch_adam(p_slice, g_slice, m, v, step, ends, hp)
            sharded_optimizer_step(flat_param, flat_grad, grad_shard, shards, adam_fn)
            assert float(flat_grad.abs().max()) == 0.0
            _torch_adam(ref_p, sum(grads)[:total], ref_m, ref_v, step, seg_ends, hparams)
            assert torch.allclose(flat_param[:total], ref_p, rtol=1e-6, atol=1e-7), (rank, step)
            assert float(flat_param[total:].abs().max()) == 0.0

        # loss accumulators: shards add up to the whole batch
        cfg = configs.small_test_config()
        bound = J.enlarged_bound(cfg)
        bb = torch.tensor(cfg['mapping']['bound'], dtype=torch.float64)
        planes, weights = J.init_planes(cfg, bound, seed=1, std=0.4), J.init_weights(cfg, seed=1)
        n = 101
        batch = synthetic.make_ray_batch(cfg, n, seed=3, zero_depth_frac=0.15)
        noise = synthetic.make_noise(n, synthetic.n_samples_total(cfg), seed=3)
        lo, hi = shard_rays(n, rank, world)
        sl = slice(lo, hi)
        with torch.no_grad():
            rend = J.render_rays(cfg, planes, weights, bound, bb, batch['rays_o'][sl], batch['rays_d'][sl], batch['target_d'][sl], noise[sl])
            acc = J.loss_accumulators(cfg, rend, batch['target_rgb'][sl], batch['target_d'][sl])
            dist.all_reduce(acc)
            got = J.losses_from_accumulators(acc, n, noise.shape[1])
            full_ret, _ = J.forward(cfg, planes, weights, bound, bb, batch['rays_o'], batch['rays_d'], batch['target_rgb'],
                                    batch['target_d'], noise)
        for k, val in got.items():
            np.testing.assert_allclose(float(val), float(full_ret[k]), rtol=1e-5, err_msg=k)
        out.put((rank, 'ok'))
    except Exception as e:      # report instead of hanging the parent
        out.put((rank, repr(e)))
    finally:
        dist.destroy_process_group()


def test_sharded_step_and_accumulators_world_size_2():
    ctx = mp.get_context('spawn')
    out = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, out)) for r in range(2)]
    for p in procs:
        p.start()
    results = [out.get(timeout=120) for _ in procs]
    for p in procs:
        p.join(timeout=60)
    assert sorted(results) == [(0, 'ok'), (1, 'ok')], results


def test_ray_sharding_covers_the_batch():
    for n in (0, 1, 7, 2048, 2049):
        for w in (1, 2, 3, 8):
            spans = [shard_rays(n, r, w) for r in range(w)]
            assert spans[0][0] == 0 and spans[-1][1] == n
            assert all(a[1] == b[0] for a, b in zip(spans, spans[1:]))
            sizes = [b - a for a, b in spans]
            assert max(sizes) - min(sizes) <= 1


def test_flat_shards_cover_segments():
    total, ends, hp = 1000, [400, 1000], ['a', 'b']
    for w in (1, 2, 3, 5):
        covered = 0
        for r in range(w):
            fs = FlatShards(total, w, r, ends, hp)
            assert fs.padded % (4 * w) == 0 and fs.local_ends[-1] == fs.shard_len
            covered += fs.shard_len
            b = 0
            for e, h in zip(fs.local_ends, fs.local_hparams):
                mid = fs.lo + (b + e) // 2
                assert h == ('a' if mid < 400 else 'b')
                b = e
        assert covered == FlatShards(total, w, 0, ends, hp).padded


def _exchange_worker(rank, world, port, out, tmpdir):
    os.environ.update({'MASTER_ADDR': '127.0.0.1', 'MASTER_PORT': str(port)})
    dist.init_process_group('gloo', rank=rank, world_size=world)
    try:
        from mneslam_b200 import JointEncoding, exchange
        cfg = configs.small_test_config(one_grid=False)
        bb = torch.tensor(cfg['mapping']['bound'], dtype=torch.float64)
        torch.manual_seed(10 + rank)
        model = JointEncoding(cfg, bb)                  # on the CPU here; the payload logic is device-agnostic
        shared = JointEncoding(cfg, bb)
        nbytes = exchange.exchange_maps(model, shared, peer=1 - rank)
        torch.manual_seed(10 + (1 - rank))
        expect = JointEncoding(cfg, bb)
        for a, b in zip(shared._flat_planes() + shared.decoder.weight_list(), expect._flat_planes() + expect.decoder.weight_list()):
            assert torch.equal(a, b)
        assert nbytes == 4 * sum(t.numel() for t in model._flat_planes() + model.decoder.weight_list())
        # the reference's file format round-trips as well
        path = os.path.join(tmpdir, 'latest_checkpoint_%d.pt' % rank)
        exchange.save_checkpoint(model, path)
        dist.barrier()
        other = exchange.load_foreign_model(JointEncoding(cfg, bb), os.path.join(tmpdir, 'latest_checkpoint_%d.pt' % (1 - rank)))
        for a, b in zip(other._flat_planes() + other.decoder.weight_list(), expect._flat_planes() + expect.decoder.weight_list()):
            assert torch.equal(a, b)
        assert sorted(exchange.checkpoint_dict(model)) == ['all_planes', 'bound', 'bounding_box', 'model']
        assert 'sdf_net.model.0.weight' in exchange.checkpoint_dict(model)['model']
        # broadcast of rank 0's map
        exchange.broadcast_map(model, src=0)
        torch.manual_seed(10)
        zero = JointEncoding(cfg, bb)
        assert all(torch.equal(a, b) for a, b in zip(model._flat_planes(), zero._flat_planes()))
        out.put((rank, 'ok'))
    except Exception as e:
        import traceback
        out.put((rank, traceback.format_exc()))
    finally:
        dist.destroy_process_group()


def test_map_exchange_between_two_agents(tmp_path):
    ctx = mp.get_context('spawn')
    out = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_exchange_worker, args=(r, 2, port, out, str(tmp_path))) for r in range(2)]
    for p in procs:
        p.start()
    results = [out.get(timeout=180) for _ in procs]
    for p in procs:
        p.join(timeout=60)
    assert sorted(results) == [(0, 'ok'), (1, 'ok')], results
