"""Map exchange between agents (BASELINE.json config 5).

In the reference every agent is an OS process pinned to one GPU; agents hand their maps to each other through
files: ``latest_checkpoint.pt`` = ``{'model': decoder state_dict, 'all_planes', 'bound', 'bounding_box'}`` written
after every mapping step (``mneslam_mp.py:294-315``) and read back by the peer into ``model_shared``
(``mp_slam/mapper.py:708-726``) -- 129 MB through the filesystem per exchange.  Here the same payload moves GPU to
GPU with ``torch.distributed`` point-to-point operations (NCCL over NVLink; gloo in the CPU tests): one flat
float32 message (planes in ``all_planes`` order, then the four decoder matrices) plus the two boxes.

The file format is kept too (``save_checkpoint`` / ``load_foreign_model``) so that checkpoints written by either
implementation load into the other.
"""
import torch
import torch.distributed as dist


def _tensors(model):
    return model._flat_planes() + model.decoder.weight_list()


def pack_map(model):
    """All map parameters as one contiguous float32 vector (logical NCHW order for the planes, so that the receiver
    does not need to know the sender's memory format)."""
    return torch.cat([t.detach().contiguous().reshape(-1) for t in _tensors(model)])


def unpack_map(model, flat):
    """Write a packed map into ``model`` (shapes must match: same config / bound)."""
    off = 0
    with torch.no_grad():
        for t in _tensors(model):
            n = t.numel()
            t.copy_(flat[off:off + n].reshape(t.shape))
            off += n
    if off != flat.numel():
        raise ValueError('packed map has %d values, model takes %d' % (flat.numel(), off))


def _boxes(model, device):
    return torch.cat([model.bound.reshape(-1).double(), model.bounding_box.reshape(-1).double().cpu()]).to(device)


def exchange_maps(model, model_shared, peer, group=None):
    """Symmetric swap with ``peer``: this agent's map goes out, the peer's map lands in ``model_shared`` (the frozen
    copy the reference keeps for distillation and loop-closure alignment).  Both agents call this."""
    mine = pack_map(model)
    theirs = torch.empty_like(mine)
    boxes_out = _boxes(model, mine.device)
    boxes_in = torch.empty_like(boxes_out)
    ops = [dist.P2POp(dist.isend, mine, peer, group), dist.P2POp(dist.irecv, theirs, peer, group),
           dist.P2POp(dist.isend, boxes_out, peer, group), dist.P2POp(dist.irecv, boxes_in, peer, group)]
    for req in dist.batch_isend_irecv(ops):
        req.wait()
    unpack_map(model_shared, theirs)
    model_shared.bound = boxes_in[:6].reshape(3, 2).float().cpu()
    model_shared.bounding_box = boxes_in[6:].reshape(3, 2).to(model_shared.bounding_box.device)
    return theirs.numel() * 4


def broadcast_map(model, src, group=None):
    """One agent's map to every rank (e.g. the fused global map after a loop closure)."""
    flat = pack_map(model)
    dist.broadcast(flat, src=src, group=group)
    if dist.get_rank(group) != src:
        unpack_map(model, flat)


def checkpoint_dict(model):
    """The reference's checkpoint payload (mneslam_mp.py:304-309)."""
    return {'model': model.state_dict(), 'all_planes': model.all_planes, 'bound': model.bound,
            'bounding_box': model.bounding_box}


def save_checkpoint(model, path):
    torch.save(checkpoint_dict(model), path)


def load_foreign_model(model_shared, path, device=None):
    """``Mapper.load_foreign_model`` (mp_slam/mapper.py:708-726): decoder weights, planes and boxes of a peer."""
    ckpt = torch.load(path, map_location=device or model_shared.device, weights_only=False)
    model_shared.load_state_dict(ckpt['model'])
    model_shared.all_planes = ckpt['all_planes']
    model_shared.bound = ckpt['bound']
    model_shared.bounding_box = ckpt['bounding_box']
    return model_shared
