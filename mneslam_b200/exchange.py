"""Map exchange between agents (BASELINE.json config 5).

In the reference every agent is an OS process pinned to one GPU; agents hand their maps to each other through
files: ``latest_checkpoint.pt`` = ``{'model': decoder state_dict, 'all_planes', 'bound', 'bounding_box'}`` written
after every mapping step (``mneslam_mp.py:294-315``) and read back by the peer into ``model_shared``
(``mp_slam/mapper.py:708-726``) -- 129 MB through the filesystem per exchange.  Here the same payload moves GPU to
GPU with ``torch.distributed`` point-to-point operations (NCCL over NVLink; gloo in the CPU tests), STRAIGHT from the
sender's parameter storage into the receiver's: one send / receive per tensor, all posted as one batch, no packing
and no staging copy.  Both sides hold the planes channels-last (a texel's 32 channels contiguous), so the bytes on the
wire are the storage itself; a plane in any other layout is re-laid out on the sending side / copied on the receiving
side (the slow path, taken only by maps that came from a reference checkpoint).

The file format is kept too (``save_checkpoint`` / ``load_foreign_model``) so that checkpoints written by either
implementation load into the other; with the hash-grid encoder (``grid.use_hash``) the tables travel as an extra key.
"""
import torch
import torch.distributed as dist


def _tensors(model):
    return model._flat_planes() + model.decoder.weight_list()


def _wire_view(t):
    """The tensor as one flat vector in WIRE order (planes: N, H, W, C -- the channels-last storage order; everything else:
    row-major), and whether that vector aliases the tensor's own storage."""
    t = t.detach()
    if t.dim() == 4:
        nhwc = t.permute(0, 2, 3, 1)
        if nhwc.is_contiguous():
            return nhwc.reshape(-1), True
        return nhwc.contiguous().reshape(-1), False
    if t.is_contiguous():
        return t.reshape(-1), True
    return t.contiguous().reshape(-1), False


def _from_wire(t, flat):
    """Write a wire-order vector into ``t`` (only needed when ``t`` could not receive in place)."""
    with torch.no_grad():
        if t.dim() == 4:
            n, c, h, w = t.shape
            t.copy_(flat.reshape(n, h, w, c).permute(0, 3, 1, 2))
        else:
            t.copy_(flat.reshape(t.shape))


def pack_map(model):
    """All map parameters as one float32 vector in wire order (a copy; the exchange itself does not pack)."""
    return torch.cat([_wire_view(t)[0] for t in _tensors(model)])


def unpack_map(model, flat):
    """Write a packed map into ``model`` (shapes must match: same config / bound)."""
    off = 0
    for t in _tensors(model):
        n = t.numel()
        _from_wire(t, flat[off:off + n])
        off += n
    if off != flat.numel():
        raise ValueError('packed map has %d values, model takes %d' % (flat.numel(), off))


def _boxes(model, device):
    return torch.cat([model.bound.reshape(-1).double(), model.bounding_box.reshape(-1).double().cpu()]).to(device)


def exchange_maps(model, model_shared, peer, group=None, recv_from=None):
    """This agent's map goes to ``peer``; the map of ``recv_from`` (default: ``peer``, a symmetric swap) lands in
    ``model_shared`` (the frozen copy the reference keeps for distillation and loop-closure alignment).  Every agent
    involved calls this; with ``peer = (rank + 1) % n`` and ``recv_from = (rank - 1) % n`` the maps travel round a ring.
    Returns the payload size in bytes (one direction)."""
    if recv_from is None:
        recv_from = peer
    outs, ins, fixups = [], [], []
    for src, dst in zip(_tensors(model), _tensors(model_shared)):
        if src.shape != dst.shape:
            raise ValueError('exchange_maps: model and model_shared differ in shape (%s vs %s)' % (tuple(src.shape), tuple(dst.shape)))
        outs.append(_wire_view(src)[0])
        view, in_place = _wire_view(dst)
        if not in_place:
            view = torch.empty_like(view)
            fixups.append((dst, view))
        ins.append(view)
    device = outs[0].device
    boxes_out = _boxes(model, device)
    boxes_in = torch.empty_like(boxes_out)
    ops = []
    for o, i in zip(outs + [boxes_out], ins + [boxes_in]):
        ops.append(dist.P2POp(dist.isend, o, peer, group))
        ops.append(dist.P2POp(dist.irecv, i, recv_from, group))
    for req in dist.batch_isend_irecv(ops):
        req.wait()
    for dst, view in fixups:
        _from_wire(dst, view)
    model_shared.bound = boxes_in[:6].reshape(3, 2).float().cpu()
    model_shared.bounding_box = boxes_in[6:].reshape(3, 2).to(model_shared.bounding_box.device)
    return 4 * sum(o.numel() for o in outs)


def broadcast_map(model, src, group=None):
    """One agent's map to every rank (e.g. the fused global map after a loop closure), tensor by tensor, in place."""
    is_src = dist.get_rank(group) == src
    works, fixups = [], []
    for t in _tensors(model):
        view, in_place = _wire_view(t)
        if not in_place and not is_src:
            fixups.append((t, view))
        works.append(dist.broadcast(view, src=src, group=group, async_op=True))
    for w in works:
        w.wait()
    for t, view in fixups:
        _from_wire(t, view)


def checkpoint_dict(model):
    """The reference's checkpoint payload (mneslam_mp.py:304-309).  With the hash-grid encoder (``grid.use_hash``, an
    extension: the reference never builds it) the tables are plain tensors outside ``state_dict()`` and ``all_planes`` is
    empty, so they are added under ``hash_tables``."""
    d = {'model': model.state_dict(), 'all_planes': model.all_planes, 'bound': model.bound,
         'bounding_box': model.bounding_box}
    if getattr(model, 'hash_meta', None) is not None:
        d['hash_tables'] = [t.detach() for t in model.hash_tables]
    return d


def save_checkpoint(model, path):
    torch.save(checkpoint_dict(model), path)


def load_foreign_model(model_shared, path, device=None):
    """``Mapper.load_foreign_model`` (mp_slam/mapper.py:708-726): decoder weights, planes (or hash tables) and boxes of
    a peer."""
    ckpt = torch.load(path, map_location=device or model_shared.device, weights_only=False)
    model_shared.load_state_dict(ckpt['model'])
    if getattr(model_shared, 'hash_meta', None) is not None:
        tables = ckpt.get('hash_tables')
        if tables is None or len(tables) != len(model_shared.hash_tables):
            raise KeyError('load_foreign_model: the model uses the hash-grid encoder but the checkpoint carries %s hash tables'
                           % ('no' if tables is None else len(tables)))
        for mine, theirs in zip(model_shared.hash_tables, tables):
            if mine.shape != theirs.shape:
                raise ValueError('load_foreign_model: hash table shape %s, checkpoint has %s' % (tuple(mine.shape), tuple(theirs.shape)))
        model_shared.hash_tables = [t.to(model_shared.hash_tables[0].device) for t in tables]
    else:
        model_shared.all_planes = ckpt['all_planes']
    model_shared.bound = ckpt['bound']
    model_shared.bounding_box = ckpt['bounding_box']
    return model_shared


class PeerMapInbox:
    """``model_shared`` re-homed into NVLink peer-mapped (symmetric) memory, so that a peer can STORE its map straight into it:
    one device-to-device copy per tensor over NVLink (``push_map``), nothing to receive, unpack or copy on this side.  All
    ranks of ``group`` construct it (collectively) for models of the same shapes.  Every tensor of ``model_shared`` keeps its
    identity, shape and strides; only its storage moves into the symmetric buffer."""

    def __init__(self, model_shared, group=None):
        import torch.distributed._symmetric_memory as symm
        group = group if group is not None else dist.group.WORLD
        self.model, self.group = model_shared, group
        tensors = _tensors(model_shared)
        self.layout, off = [], 0
        for t in tensors:
            if not _wire_view(t)[1]:
                raise ValueError('PeerMapInbox: model_shared must hold dense (channels-last / contiguous) tensors')
            self.layout.append((tuple(t.shape), t.stride(), off, t.numel()))
            off += t.numel()
        self.total = off + 24                          # + the two boxes (12 doubles)
        device = tensors[0].device
        self.buf = symm.empty(self.total, dtype=torch.float32, device=device)
        self.handle = symm.rendezvous(self.buf, group.group_name)
        with torch.no_grad():
            for t, (shape, stride, o, n) in zip(tensors, self.layout):
                view = torch.as_strided(self.buf, shape, stride, o)
                view.copy_(t)
                t.data = view
        self.rank = dist.get_rank(group)

    def flat_on(self, rank):
        """The whole inbox of ``rank`` as a flat float32 tensor (peer-mapped when ``rank`` is not this rank)."""
        return self.handle.get_buffer(rank, (self.total,), torch.float32)

    def adopt_boxes(self):
        """After a push (and the barrier): read the sender's two boxes out of the inbox."""
        boxes = self.buf[self.total - 24:].view(torch.float64).clone()
        self.model.bound = boxes[:6].reshape(3, 2).float().cpu()
        self.model.bounding_box = boxes[6:].reshape(3, 2).to(self.model.bounding_box.device)


def push_map(model, inbox, peer, barrier=True):
    """Store this agent's map into ``peer``'s ``PeerMapInbox`` (= its ``model_shared``): one NVLink device-to-device copy per
    tensor from the parameter storage, then a cross-rank barrier (every rank of the inbox's group calls this, e.g. round a
    ring) after which each rank's ``model_shared`` holds the map that was pushed to it.  Returns the bytes sent."""
    dst = inbox.flat_on(peer)
    n_bytes = 0
    with torch.no_grad():
        for t, (shape, stride, off, n) in zip(_tensors(model), inbox.layout):
            if tuple(t.shape) != shape:
                raise ValueError('push_map: tensor shape %s does not match the inbox (%s)' % (tuple(t.shape), shape))
            dst[off:off + n].copy_(_wire_view(t)[0])
            n_bytes += 4 * n
        boxes = _boxes(model, dst.device)
        dst[inbox.total - 24:].view(torch.float64).copy_(boxes)
    if barrier:
        inbox.handle.barrier(channel=0)
        inbox.adopt_boxes()
    return n_bytes
