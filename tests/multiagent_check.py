"""BASELINE.json configs[4] in small: three agents, one per GPU (run under torchrun with 3 ranks), each mapping its own
synthetic Replica-shaped trajectory on its own map; then one map exchange round a ring over NVLink (the reference writes
``latest_checkpoint.pt`` to disk and the peer reads it back, mneslam_mp.py:294-315 / mp_slam/mapper.py:708-726) and ten
iterations of the joint distillation (mp_slam/mapper.py:594-644) with the received map as the teacher.

Checks: the received map is bit-identical to what the sender holds (checksums gathered over the ranks), the distillation
loss is finite and falls.  Prints the exchange time against the file round trip it replaces."""
import os
import sys
import tempfile
import time

import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

from mneslam_b200 import JointEncoding, configs, exchange, synthetic      # noqa: E402
from mneslam_b200.mapping import DeviceKeyframeRays, FusedMapper          # noqa: E402


def checksum(model):
    return torch.stack([t.detach().double().sum() for t in model._flat_planes() + model.decoder.weight_list()])


def main():
    rank, world, local = int(os.environ['RANK']), int(os.environ['WORLD_SIZE']), int(os.environ['LOCAL_RANK'])
    torch.cuda.set_device(local)
    dev = torch.device('cuda', local)
    dist.init_process_group('nccl', device_id=dev)
    cfg = configs.replica_room0()
    bb = torch.tensor(cfg['mapping']['bound'], dtype=torch.float64)
    torch.manual_seed(100 + rank)                              # every agent starts from its own map
    model = JointEncoding(cfg, bb).to(dev)
    model.train()
    torch.manual_seed(0)
    shared = JointEncoding(cfg, bb).to(dev)                    # model_shared: the frozen copy of a peer's map
    mapper = FusedMapper(model)                                # one agent = one GPU = one map (world_size 1)
    rays, poses = synthetic.make_keyframe_db(cfg, n_kf=6, seed=10 * rank)      # its own trajectory
    db = DeviceKeyframeRays(rays.to(dev), poses.to(dev))
    for _ in range(20):
        mapper.step(*db.sample(2048))
    torch.cuda.synchronize()
    dist.barrier()

    # ---- the exchange: round a ring, straight from parameter storage to parameter storage
    nxt, prv = (rank + 1) % world, (rank - 1) % world
    exchange.exchange_maps(model, shared, nxt, recv_from=prv)              # warm-up (NCCL connections)
    torch.cuda.synchronize(); dist.barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    times = []
    for _ in range(5):
        dist.barrier(); torch.cuda.synchronize()
        e0.record()
        nbytes = exchange.exchange_maps(model, shared, nxt, recv_from=prv)
        e1.record()
        torch.cuda.synchronize()
        times.append(e0.elapsed_time(e1))
    ms = torch.tensor([sorted(times)[2]], device=dev, dtype=torch.float64)
    dist.all_reduce(ms, op=dist.ReduceOp.MAX)
    sums = [torch.zeros_like(checksum(model)) for _ in range(world)]
    dist.all_gather(sums, checksum(model))
    assert torch.equal(checksum(shared), sums[prv]), 'rank %d: the received map differs from what rank %d holds' % (rank, prv)
    assert torch.allclose(shared.bound, model.bound) and torch.allclose(shared.bounding_box.cpu(), model.bounding_box.cpu())

    # ---- the same exchange as peer stores into model_shared itself (PeerMapInbox: model_shared lives in NVLink peer memory)
    torch.manual_seed(0)
    shared2 = JointEncoding(cfg, bb).to(dev)
    inbox = exchange.PeerMapInbox(shared2)
    exchange.push_map(model, inbox, nxt)
    torch.cuda.synchronize(); dist.barrier()
    ptimes = []
    for _ in range(5):
        dist.barrier(); torch.cuda.synchronize()
        e0.record()
        exchange.push_map(model, inbox, nxt)
        e1.record()
        torch.cuda.synchronize()
        ptimes.append(e0.elapsed_time(e1))
    pms = torch.tensor([sorted(ptimes)[2]], device=dev, dtype=torch.float64)
    dist.all_reduce(pms, op=dist.ReduceOp.MAX)
    assert torch.equal(checksum(shared2), sums[prv]), 'rank %d: the pushed map differs from what rank %d holds' % (rank, prv)

    # ---- the file round trip the reference does instead (rank 0 only; local temporary directory)
    file_ms = None
    if rank == 0:
        with tempfile.TemporaryDirectory() as tmp:
            path = os.path.join(tmp, 'latest_checkpoint.pt')
            torch.cuda.synchronize()
            t0 = time.perf_counter()
            exchange.save_checkpoint(model, path)
            exchange.load_foreign_model(shared, path, device=dev)
            torch.cuda.synchronize()
            file_ms = (time.perf_counter() - t0) * 1e3
    dist.barrier()
    exchange.exchange_maps(model, shared, nxt, recv_from=prv)              # (rank 0's shared copy was overwritten by the file test)

    # ---- joint distillation: the received map teaches this agent's map on the rays of a shared keyframe
    shared.eval()
    ro, rd, _, _ = db.sample(1024)
    first = last = None
    for it in range(10):
        losses = mapper.distill_step(shared, ro, rd)
        # (a teacher that renders a depth below the truncation distance leaves the free-space mask empty: that loss is then a
        #  mean over nothing, nan, as in the reference -- it contributes no gradient; the check follows the other terms)
        val = float(torch.nansum(losses[:7] * mapper.loss_weights[:7]))
        first = val if first is None else first
        last = val
    assert first == first and last == last and last < first, (first, last)
    torch.cuda.synchronize()
    dist.barrier()
    if rank == 0:
        print('multiagent_check ok: %d agents, map %.1f MB, exchange %.3f ms by NCCL send/recv (%.0f GB/s per direction), %.3f ms by '
              'peer stores into model_shared (%.0f GB/s; max over ranks), checkpoint file round trip %.0f ms; distillation loss %.4f -> %.4f'
              % (world, nbytes / 1e6, float(ms), nbytes / float(ms) / 1e6, float(pms), nbytes / float(pms) / 1e6, file_ms, first, last), flush=True)
    dist.barrier()
    os._exit(0)


if __name__ == '__main__':
    main()
