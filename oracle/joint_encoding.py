"""CPU oracle: restatement of the reference's JointEncoding ray-batch path in plain torch (float32, CPU).

TEST INFRASTRUCTURE ONLY (see ``oracle/__init__.py``).  Every function names the reference lines it
follows; ``oracle/make_golden.py`` pins it against the reference's own modules.

State is passed explicitly:
  ``planes``  tuple of 3 (one grid) or 6 (two grids) lists of ``[1, C, H, W]`` tensors -- the layout of
              ``JointEncoding.all_planes`` (``model/scene_rep.py:85-117,163-168``);
  ``weights`` dict ``sdf0 [32,112] sdf1 [16,32] col0 [32,63|127] col1 [3,32]`` -- the four bias-free
              ``nn.Linear`` matrices of ``ColorSDFNet(_v2)`` (``model/decoder.py:39-55,92-108``);
  ``bound``   float32 ``[3,2]`` enlarged box (``model/scene_rep.py:80-83``), ``bounding_box`` float64
              ``[3,2]`` raw box (``mneslam_mp.py:223``).
"""
import math

import torch
import torch.nn.functional as F

from .oneblob import oneblob


# --------------------------------------------------------------------------------------------
# construction helpers
# --------------------------------------------------------------------------------------------
def enlarged_bound(cfg):
    """``load_bound`` (model/scene_rep.py:72-83): grow the upper corner to a multiple of
    ``bound_dividable`` -- float32 arithmetic, truncation through ``.int()``."""
    b = torch.tensor(cfg['mapping']['bound'], dtype=torch.float64) * cfg['scale']
    b = b.float()
    q = cfg['planes_res']['bound_dividable']
    b[:, 1] = (((b[:, 1] - b[:, 0]) / q).int() + 1) * q + b[:, 0]
    return b


def plane_shapes(cfg, bound, colour=False):
    """Shapes of the xy / xz / yz planes at (coarse, fine) resolution
    (model/scene_rep.py:92-109,126-142): ``int(len / res)`` in float32, xy=[Ny,Nx] xz=[Nz,Nx] yz=[Nz,Ny]."""
    key = 'c_planes_res' if colour else 'planes_res'
    ext = bound[:, 1] - bound[:, 0]
    out = []
    for res in (cfg[key]['coarse'], cfg[key]['fine']):
        nx, ny, nz = [int(v) for v in (ext / res).tolist()]
        out.append({'xy': (ny, nx), 'xz': (nz, nx), 'yz': (nz, ny)})
    return out


def init_planes(cfg, bound, seed=0, std=0.01):
    gen = torch.Generator().manual_seed(seed)
    c = cfg['model']['c_dim']
    groups = [False] if cfg['grid']['oneGrid'] else [False, True]
    planes = []
    for colour in groups:
        shapes = plane_shapes(cfg, bound, colour)
        for axis in ('xy', 'xz', 'yz'):
            planes.append([torch.randn(1, c, *s[axis], generator=gen) * std for s in shapes])
    return tuple(planes)


def init_weights(cfg, seed=0):
    gen = torch.Generator().manual_seed(seed + 1)
    d, m = cfg['decoder'], cfg['model']
    feat = 32 if cfg['grid'].get('use_hash', False) else m['input_ch']
    in_sdf = feat + m['input_ch_pos']
    in_col = m['input_ch_pos'] + d['geo_feat_dim'] + (0 if cfg['grid']['oneGrid'] else feat)

    def lin(o, i):   # nn.Linear default init: U(-1/sqrt(i), 1/sqrt(i))
        k = 1.0 / math.sqrt(i)
        return (torch.rand(o, i, generator=gen) * 2 - 1) * k
    return {'sdf0': lin(d['hidden_dim'], in_sdf), 'sdf1': lin(1 + d['geo_feat_dim'], d['hidden_dim']),
            'col0': lin(d['hidden_dim_color'], in_col), 'col1': lin(3, d['hidden_dim_color'])}


# --------------------------------------------------------------------------------------------
# sampling  (model/scene_rep.py:361-381)
# --------------------------------------------------------------------------------------------
def sample_tables(cfg):
    """The four constant linspace tables the reference builds on the CPU every call."""
    t, cam = cfg['training'], cfg['cam']
    return {
        'uniform_d': torch.linspace(cam['near'], cam['far'], t['n_samples_d']) if t['n_samples_d'] > 0
        else torch.empty(0),
        'range_off': torch.linspace(-t['range_d'], t['range_d'], steps=t['n_range_d']),
        'range_fallback': torch.linspace(cam['near'], cam['far'], steps=t['n_range_d']),
        'uniform': torch.linspace(cam['near'], cam['far'], t.get('n_samples', 0) or 1),
    }


def sample_z(cfg, n_rays, target_d=None, noise=None):
    """Depth-guided + uniform samples merged by sort, then stratified perturbation.
    Returns (z_vals [N,S], z_sorted [N,S] before perturbation)."""
    tab = sample_tables(cfg)
    dev = target_d.device if target_d is not None else (noise.device if noise is not None else torch.device('cpu'))
    tab = {k: v.to(dev) for k, v in tab.items()}          # built on the CPU, moved -- as the reference does (:362-374)
    if target_d is not None:
        z_near = tab['range_off'][None, :].repeat(n_rays, 1) + target_d                 # :363-364
        z_near[target_d.squeeze(-1) <= 0] = tab['range_fallback']                       # :365
        if cfg['training']['n_samples_d'] > 0:
            z = torch.cat([tab['uniform_d'][None, :].repeat(n_rays, 1), z_near], -1)    # :368-369
            z, _ = torch.sort(z, -1)
        else:
            z = z_near
    else:
        z = tab['uniform'][None, :].repeat(n_rays, 1)                                   # :373-374
    z_sorted = z
    if cfg['training']['perturb'] > 0.:                                                 # :377-381
        mid = 0.5 * (z[:, 1:] + z[:, :-1])
        hi = torch.cat([mid, z[:, -1:]], -1)
        lo = torch.cat([z[:, :1], mid], -1)
        z = lo + (hi - lo) * noise
    return z, z_sorted


def sample_pdf(bins, weights, n_importance, det, u=None):
    """Inverse-CDF importance sampling (model/utils.py:65-104).  Returns (samples, inds) -- ``inds`` are
    the int64 searchsorted bin indices that must match bit for bit."""
    w = weights + 1e-5
    pdf = w / w.sum(-1, keepdim=True)
    cdf = torch.cat([torch.zeros_like(pdf[:, :1]), torch.cumsum(pdf, -1)], -1)
    if det:
        u = torch.linspace(0.5 / n_importance, 1. - 0.5 / n_importance, steps=n_importance)
        u = u.expand(cdf.shape[0], n_importance)
    u = u.contiguous()
    inds = torch.searchsorted(cdf, u, right=True)
    lo = (inds - 1).clamp(min=0)
    hi = inds.clamp(max=cdf.shape[-1] - 1)
    c_lo, c_hi = torch.gather(cdf, 1, lo), torch.gather(cdf, 1, hi)
    b_lo, b_hi = torch.gather(bins, 1, lo), torch.gather(bins, 1, hi)
    den = c_hi - c_lo
    den = torch.where(den < 1e-5, torch.ones_like(den), den)
    return b_lo + (u - c_lo) / den * (b_hi - b_lo), inds


# --------------------------------------------------------------------------------------------
# encoders
# --------------------------------------------------------------------------------------------
def normalize_plane_coords(p, bound):
    """model/utils.py:27-41: per axis ``((p - lo) / (hi - lo)) * 2 - 1`` in float32."""
    cols = [((p[:, a] - bound[a, 0]) / (bound[a, 1] - bound[a, 0])) * 2 - 1.0 for a in range(3)]
    return torch.stack(cols, -1)


def bilinear_border(plane, gx, gy):
    """Index-arithmetic restatement of ``F.grid_sample(mode='bilinear', padding_mode='border',
    align_corners=True)`` for one plane ``[1,C,H,W]`` at normalised coords (gx -> W axis, gy -> H axis)
    (model/scene_rep.py:43-48; SURVEY.md 8(a) note 1).  Gradient w.r.t. a coordinate is zero where its
    pixel coordinate was clamped (ATen ``clip_coordinates_set_grad``)."""
    _, C, H, W = plane.shape
    tex = plane[0].permute(1, 2, 0)                       # [H, W, C]

    def axis(g, size):
        u = ((g + 1) / 2) * (size - 1)
        inside = (u > 0) & (u < size - 1)
        u = torch.where(inside, u, u.detach().clamp(0, size - 1))
        i0 = torch.floor(u.detach())
        f = u - i0
        i0 = i0.long()
        i1 = i0 + 1
        ok1 = (i1 <= size - 1).to(u.dtype)
        return i0, i1.clamp(max=size - 1), 1 - f, f * ok1

    x0, x1, wx0, wx1 = axis(gx, W)
    y0, y1, wy0, wy1 = axis(gy, H)
    return (tex[y0, x0] * (wy0 * wx0)[:, None] + tex[y0, x1] * (wy0 * wx1)[:, None]
            + tex[y1, x0] * (wy1 * wx0)[:, None] + tex[y1, x1] * (wy1 * wx1)[:, None])


def triplane_features(p_nor, planes_xy, planes_xz, planes_yz, fast=False):
    """``sample_plane_feature`` (model/scene_rep.py:28-53): per level xy+xz+yz, levels concatenated.
    ``fast=True`` calls torch's own grid_sample kernel (what the reference executes -- used for the timed
    CPU baseline); ``fast=False`` uses the explicit restatement above."""
    out = []
    for lvl in range(len(planes_xy)):
        if fast:
            g = p_nor[None, :, None]

            def samp(pl, a, b):
                return F.grid_sample(pl, g[..., [a, b]], padding_mode='border', align_corners=True,
                                     mode='bilinear')[0, :, :, 0].t()
            xy, xz, yz = samp(planes_xy[lvl], 0, 1), samp(planes_xz[lvl], 0, 2), samp(planes_yz[lvl], 1, 2)
        else:
            xy = bilinear_border(planes_xy[lvl], p_nor[:, 0], p_nor[:, 1])
            xz = bilinear_border(planes_xz[lvl], p_nor[:, 0], p_nor[:, 2])
            yz = bilinear_border(planes_yz[lvl], p_nor[:, 1], p_nor[:, 2])
        out.append(xy + xz + yz)
    return torch.cat(out, -1)


def blob_inputs(pts, bounding_box):
    """model/scene_rep.py:292: normalisation by the RAW box in float64, cast to float32 by tcnn."""
    bb = bounding_box.double()
    return ((pts.double() - bb[:, 0]) / (bb[:, 1] - bb[:, 0])).float()


# --------------------------------------------------------------------------------------------
# decoder (model/decoder.py:110-175) and point query (model/scene_rep.py:273-301)
# --------------------------------------------------------------------------------------------
def decode(weights, feat, blob, cfeat=None):
    h = torch.relu(torch.cat([feat, blob], -1) @ weights['sdf0'].t()) @ weights['sdf1'].t()
    sdf, geo = h[:, :1], h[:, 1:]
    cin = [blob, geo] if cfeat is None else [blob, cfeat, geo]
    rgb = torch.relu(torch.cat(cin, -1) @ weights['col0'].t()) @ weights['col1'].t()
    return torch.cat([rgb, sdf], -1), geo


def query_points(cfg, planes, weights, bound, bounding_box, pts, fast=False):
    """pts [M,3] -> raw [M,4] = (rgb logits, sdf).  With ``cfg['grid']['use_hash']`` the parametric encoder is the
    multi-resolution hash grid (``oracle/hashgrid.py``) and ``planes`` is ``([table],)`` or ``([table], [colour table])``."""
    if cfg['grid'].get('use_hash', False):
        from . import hashgrid
        x01 = blob_inputs(pts, bounding_box)
        feat = hashgrid.encode(hashgrid.make_meta(cfg), planes[0][0], x01)
        cfeat = None
        if not cfg['grid']['oneGrid']:      # the colour encoder of the (commented) model/scene_rep.py:169: resolution from voxel_color
            cmeta = hashgrid.make_meta(cfg, resolution=hashgrid.desired_resolution(cfg, 'voxel_color'))
            cfeat = hashgrid.encode(cmeta, planes[1][0], x01)
        return decode(weights, feat, oneblob(x01, cfg['pos']['n_bins']), cfeat)[0]
    p_nor = normalize_plane_coords(pts, bound)
    feat = triplane_features(p_nor, planes[0], planes[1], planes[2], fast)
    blob = oneblob(blob_inputs(pts, bounding_box), cfg['pos']['n_bins'])
    cfeat = None
    if not cfg['grid']['oneGrid']:
        cfeat = triplane_features(p_nor, planes[3], planes[4], planes[5], fast)
    return decode(weights, feat, blob, cfeat)[0]


# --------------------------------------------------------------------------------------------
# compositing (model/scene_rep.py:183-230)
# --------------------------------------------------------------------------------------------
def sdf_to_weights(cfg, sdf, z):
    tr = cfg['training']['trunc']
    w = torch.sigmoid(sdf / tr) * torch.sigmoid(-sdf / tr)
    crossing = (sdf[:, 1:] * sdf[:, :-1] < 0.0).float()
    inds = torch.argmax(crossing, dim=1)                     # 0 when the ray has no sign change
    z_min = torch.gather(z, 1, inds[:, None])
    keep = (z < z_min + cfg['data']['sc_factor'] * tr).float()
    w = w * keep
    return w / (w.sum(-1, keepdim=True) + 1e-8), inds


def composite(cfg, raw, z):
    rgb = torch.sigmoid(raw[..., :3])
    w, inds = sdf_to_weights(cfg, raw[..., 3], z)
    rgb_map = (w[..., None] * rgb).sum(-2)
    depth = (w * z).sum(-1)
    var = (w * (z - depth[:, None]) ** 2).sum(-1)
    acc = w.sum(-1)
    disp = 1. / torch.max(torch.full_like(depth, 1e-10), depth / acc)
    if cfg['training']['white_bkgd']:
        rgb_map = rgb_map + (1. - acc[:, None])
    return {'rgb': rgb_map, 'depth': depth, 'disp_map': disp, 'acc_map': acc, 'depth_var': var,
            'weights': w, 'inds': inds}


def render_rays(cfg, planes, weights, bound, bounding_box, rays_o, rays_d, target_d=None, noise=None,
                u_importance=None, fast=False):
    """model/scene_rep.py:351-419."""
    n = rays_o.shape[0]
    z, z_sorted = sample_z(cfg, n, target_d, noise)

    def shade(zv):
        pts = rays_o[:, None, :] + rays_d[:, None, :] * zv[:, :, None]
        raw = query_points(cfg, planes, weights, bound, bounding_box, pts.reshape(-1, 3), fast)
        raw = raw.reshape(n, zv.shape[1], 4)
        return raw, composite(cfg, raw, zv)

    raw, out = shade(z)
    ret = {}
    n_imp = cfg['training']['n_importance']
    if n_imp > 0:                                                                       # :389-401
        first = out
        mid = 0.5 * (z[:, 1:] + z[:, :-1])
        z_new, pdf_inds = sample_pdf(mid, first['weights'][:, 1:-1].detach(), n_imp,
                                     det=(cfg['training']['perturb'] == 0.), u=u_importance)
        z_new = z_new.detach()
        z, _ = torch.sort(torch.cat([z, z_new], -1), -1)
        raw, out = shade(z)
        ret.update({'rgb0': first['rgb'], 'depth0': first['depth'], 'disp0': first['disp_map'],
                    'acc0': first['acc_map'], 'depth_var0': first['depth_var'],
                    'z_std': torch.std(z_new, dim=-1, unbiased=False), 'pdf_inds': pdf_inds,
                    'weights0': first['weights']})
    ret.update(out)
    ret.update({'z_vals': z, 'raw': raw, 'z_sorted': z_sorted})
    return ret


# --------------------------------------------------------------------------------------------
# losses (model/scene_rep.py:475-511,549-611; model/utils.py:117-185)
# --------------------------------------------------------------------------------------------
def _masked_mean(x, mask):
    m = mask.to(x.dtype)
    return (x * m).sum() / m.sum()          # 0/0 = nan on an empty mask, like mean() of an empty select


def losses(cfg, rend, target_rgb, target_d):
    d = target_d.squeeze(-1)
    z, sdf = rend['z_vals'], rend['raw'][..., 3]
    valid = (d > 0.) & (d < cfg['cam']['depth_trunc'])
    rgb_loss = ((rend['rgb'] - target_rgb) ** 2).mean()
    psnr = -10. * torch.log(rgb_loss) / math.log(10.)       # psnr is taken before the rgb0 term is added
    depth_loss = _masked_mean((rend['depth'] - d) ** 2, valid)
    if 'rgb0' in rend:
        rgb_loss = rgb_loss + ((rend['rgb0'] - target_rgb) ** 2).mean()
        depth_loss = depth_loss + _masked_mean((rend['depth0'] - d) ** 2, valid)

    # Co-SLAM pair: masks over every ray, means over all N*S samples, balanced by sample counts
    T = cfg['training']['trunc'] * cfg['data']['sc_factor']
    front = (z < (d - T)[:, None]).float()
    back = (z > (d + T)[:, None]).float()
    near = (1. - front) * (1. - back) * (d > 0.).float()[:, None]
    n_fs, n_sdf = torch.count_nonzero(front), torch.count_nonzero(near)
    fs_w = 1.0 - n_fs / (n_fs + n_sdf)
    sdf_w = 1.0 - n_sdf / (n_fs + n_sdf)
    co_fs = ((sdf * front - front) ** 2).mean() * fs_w
    co_sdf = (((z + sdf * T) * near - d[:, None] * near) ** 2).mean() * sdf_w

    # ESLAM triple: rays with a positive depth only, each term the mean over its own mask
    Tm = cfg['model']['truncation']
    has_d = (d > 0)[:, None]
    e_front = (z < (d - Tm)[:, None]) & has_d
    e_back = (z > (d + Tm)[:, None]) & has_d
    e_center = (z > (d - 0.4 * Tm)[:, None]) & (z < (d + 0.4 * Tm)[:, None]) & has_d
    e_tail = (~e_front) & (~e_back) & (~e_center) & has_d
    resid = (z + sdf * Tm - d[:, None]) ** 2
    return {'rgb': rend['rgb'], 'depth': rend['depth'], 'rgb_loss': rgb_loss, 'depth_loss': depth_loss,
            'co_sdf_loss': co_sdf, 'co_fs_loss': co_fs,
            'e_fs_loss': _masked_mean((sdf - 1.0) ** 2, e_front),
            'e_center_loss': _masked_mean(resid, e_center),
            'e_tail_loss': _masked_mean(resid, e_tail), 'psnr': psnr.reshape(1)}


def total_loss(cfg, ret, is_co_sdf=None):
    """``MNESLAM.get_loss_from_ret`` (mneslam_mp.py:350-372) with rgb, depth and sdf terms on."""
    t, m = cfg['training'], cfg['mapping']
    if is_co_sdf is None:
        is_co_sdf = cfg['is_co_sdf']
    loss = t['rgb_weight'] * ret['rgb_loss'] + t['depth_weight'] * ret['depth_loss']
    if is_co_sdf:
        return loss + t['sdf_weight'] * ret['co_sdf_loss'] + t['fs_weight'] * ret['co_fs_loss']
    return (loss + m['w_sdf_fs'] * ret['e_fs_loss'] + m['w_sdf_center'] * ret['e_center_loss']
            + m['w_sdf_tail'] * ret['e_tail_loss'])


def forward(cfg, planes, weights, bound, bounding_box, rays_o, rays_d, target_rgb, target_d, noise,
            u_importance=None, fast=False):
    """``JointEncoding.forward`` in training mode (model/scene_rep.py:549-611) -> (loss dict, render dict)."""
    rend = render_rays(cfg, planes, weights, bound, bounding_box, rays_o, rays_d, target_d, noise,
                       u_importance, fast)
    return losses(cfg, rend, target_rgb, target_d), rend


# --------------------------------------------------------------------------------------------
# pose parameterisations that consume the ray gradients (optimization/utils.py:161-210,
# optimization/pose_array.py:6-29)
# --------------------------------------------------------------------------------------------
def axis_angle_to_matrix(r):
    theta = r.norm(dim=-1, keepdim=True)
    k = r / theta
    z = torch.zeros_like(k[..., 0])
    K = torch.stack([torch.stack([z, -k[..., 2], k[..., 1]], -1),
                     torch.stack([k[..., 2], z, -k[..., 0]], -1),
                     torch.stack([-k[..., 1], k[..., 0], z], -1)], -2)
    eye = torch.eye(3, dtype=r.dtype, device=r.device).expand(K.shape)
    return eye + torch.sin(theta)[..., None] * K + (1. - torch.cos(theta))[..., None] * (K @ K)


def quat_to_matrix(q):
    """Real-first quaternion, not assumed normalised (pytorch3d ``quaternion_to_matrix`` /
    ``quad2rotation``)."""
    r, i, j, k = q.unbind(-1)
    s = 2.0 / (q * q).sum(-1)
    m = torch.stack([1 - s * (j * j + k * k), s * (i * j - k * r), s * (i * k + j * r),
                     s * (i * j + k * r), 1 - s * (i * i + k * k), s * (j * k - i * r),
                     s * (i * k - j * r), s * (j * k + i * r), 1 - s * (i * i + j * j)], -1)
    return m.reshape(q.shape[:-1] + (3, 3))


def rays_from_pose(R, t, dirs_cam):
    """mp_slam/mapper.py:392-393: ``rays_d = sum_k dirs_k R[:, k]``, ``rays_o = t``."""
    return t[None, :].expand(dirs_cam.shape[0], 3), torch.sum(dirs_cam[:, None, :] * R, -1)


# --------------------------------------------------------------------------------------------
# the 16 loss accumulators of include/mneslam_b200.h, restated for the multi-GPU tests: every slot is a plain
# sum over rays / samples, so shards add up; the losses are ratios of the summed slots.
# --------------------------------------------------------------------------------------------
def loss_accumulators(cfg, rend, target_rgb, target_d):
    d = target_d.squeeze(-1)
    z, sdf = rend['z_vals'], rend['raw'][..., 3]
    valid = (d > 0.) & (d < cfg['cam']['depth_trunc'])
    T = cfg['training']['trunc'] * cfg['data']['sc_factor']
    front = (z < (d - T)[:, None])
    back = (z > (d + T)[:, None])
    near = (~front) & (~back) & (d > 0.)[:, None]
    Tm = cfg['model']['truncation']
    has_d = (d > 0)[:, None]
    e_front = (z < (d - Tm)[:, None]) & has_d
    e_back = (z > (d + Tm)[:, None]) & has_d
    e_center = (z > (d - 0.4 * Tm)[:, None]) & (z < (d + 0.4 * Tm)[:, None]) & has_d
    e_tail = (~e_front) & (~e_back) & (~e_center) & has_d
    resid_e = (z + sdf * Tm - d[:, None]) ** 2
    resid_co = (z + sdf * T - d[:, None]) ** 2
    fs = (sdf - 1.0) ** 2
    acc = torch.zeros(16, dtype=torch.float64)
    acc[0], acc[1], acc[2] = valid.sum(), front.sum(), near.sum()
    acc[3], acc[4], acc[5] = e_front.sum(), e_center.sum(), e_tail.sum()
    acc[6] = ((rend['rgb'] - target_rgb) ** 2).double().sum()
    acc[7] = (((rend['depth'] - d) ** 2) * valid).double().sum()
    acc[8], acc[9] = (fs * front).double().sum(), (resid_co * near).double().sum()
    acc[10], acc[11], acc[12] = (fs * e_front).double().sum(), (resid_e * e_center).double().sum(), (resid_e * e_tail).double().sum()
    return acc


def losses_from_accumulators(acc, n_rays_total, n_samples):
    N, NS = float(n_rays_total), float(n_rays_total) * n_samples
    fs_w = 1.0 - acc[1] / (acc[1] + acc[2])
    sdf_w = 1.0 - acc[2] / (acc[1] + acc[2])
    return {'rgb_loss': acc[6] / (3 * N), 'depth_loss': acc[7] / acc[0], 'co_fs_loss': acc[8] / NS * fs_w,
            'co_sdf_loss': acc[9] / NS * sdf_w, 'e_fs_loss': acc[10] / acc[3], 'e_center_loss': acc[11] / acc[4],
            'e_tail_loss': acc[12] / acc[5]}
