"""Golden fixtures for the inference entry points and the pose parameterisation, from the REFERENCE'S OWN code.

Run in the build container only (``/root/reference`` does not exist on the GPU box):

    python -m oracle.make_golden_inference

Writes ``tests/golden/small_inference.npz``:

* ``render_img`` (model/scene_rep.py:422-473) on a small camera, with ``gt_depth`` given and ``None`` -- this pins
  ``get_rays`` (model/utils.py:6-25) and the raw ``cam.cx / cam.cy`` convention of ``:438-441``;
* ``render_surface_color`` (model/scene_rep.py:333-349);
* eval-mode ``forward`` (model/scene_rep.py:566-567: the render dict instead of the losses);
* ``axis_angle_to_matrix`` / ``at_to_transform_matrix`` (optimization/utils.py:161-197): the module imports pytorch3d at its
  top (absent here), so the two function definitions are cut out of the reference's source file by line range and executed
  on their own -- they need torch only.  (``qt_to_transform_matrix`` calls pytorch3d's ``quaternion_to_matrix``: un-vendored
  third party, stays pinned by restatement only.)

``training.perturb`` is 0 in this fixture: the reference renders an image in chunks of ``ray_batch_size`` rays and draws
its stratified noise per chunk, so a perturbed image is not reproducible by an implementation that chunks differently.

TEST INFRASTRUCTURE ONLY.
"""
import os
import sys

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
REF = '/root/reference'
sys.path.insert(0, ROOT)

from oracle import joint_encoding as J                       # noqa: E402
from oracle.make_golden import reference_model               # noqa: E402
from mneslam_b200 import configs, synthetic                  # noqa: E402


def inference_config():
    cfg = configs.small_test_config(one_grid=True)
    cfg['training']['perturb'] = 0
    cfg['cam'].update({'H': 28, 'W': 36, 'fx': 30.0, 'fy': 31.0, 'cx': 17.5, 'cy': 13.5, 'crop_edge': 2})
    return cfg


def reference_pose_functions():
    """axis_angle_to_matrix and at_to_transform_matrix, executed from the reference's own source lines."""
    path = os.path.join(REF, 'optimization', 'utils.py')
    lines = open(path).read().split('\n')

    def cut(name):
        start = next(i for i, l in enumerate(lines) if l.startswith('def %s(' % name))
        end = next(i for i in range(start + 1, len(lines)) if lines[i].startswith('def ') or lines[i].startswith('# TODO'))
        return '\n'.join(lines[start:end])
    ns = {'torch': torch}
    exec(cut('axis_angle_to_matrix'), ns)
    exec(cut('at_to_transform_matrix'), ns)
    return ns['axis_angle_to_matrix'], ns['at_to_transform_matrix']


def main():
    os.chdir(REF)
    cfg = inference_config()
    seed = 12
    bound = J.enlarged_bound(cfg)
    planes = J.init_planes(cfg, bound, seed=seed, std=0.4)
    weights = J.init_weights(cfg, seed=seed)
    model, _, bb = reference_model(cfg, planes, weights)
    model.ray_batch_size = 100                       # several ragged chunks
    cam = cfg['cam']
    H, W = cam['H'] - 2 * cam['crop_edge'], cam['W'] - 2 * cam['crop_edge']
    gen = torch.Generator().manual_seed(seed)
    c2w = synthetic.make_pose(cfg, gen)
    gt_depth = torch.rand(H, W, generator=gen) * 2.0 + 0.3
    gt_depth[torch.rand(H, W, generator=gen) < 0.1] = 0.0
    out = {'seed': seed, 'c2w': c2w.numpy(), 'gt_depth': gt_depth.numpy(), 'H': H, 'W': W}
    model.eval()
    d0, c0 = model.render_img(c2w, 'cpu', gt_depth=None)
    d1, c1 = model.render_img(c2w, 'cpu', gt_depth=gt_depth.clone())
    assert d0.dtype == torch.float64 and tuple(d0.shape) == (H, W) and tuple(c0.shape) == (H, W, 3)
    out.update(img_depth_nodepth=d0.numpy(), img_color_nodepth=c0.numpy(), img_depth=d1.numpy(), img_color=c1.numpy())

    # render_surface_color: points [N,1,3] on a sphere-ish cloud inside the box, normals random unit vectors
    n = 57
    b = torch.tensor(cfg['mapping']['bound'], dtype=torch.float32)
    pts = b[:, 0] + (b[:, 1] - b[:, 0]) * (0.2 + 0.6 * torch.rand(n, 3, generator=gen))
    normal = torch.nn.functional.normalize(torch.randn(n, 3, generator=gen), dim=-1)
    with torch.no_grad():
        rgb = model.render_surface_color(pts[:, None, :], normal)
    out.update(surf_pts=pts.numpy(), surf_normal=normal.numpy(), surf_rgb=rgb.numpy())

    # eval-mode forward
    batch = synthetic.make_ray_batch(cfg, 41, seed=seed)
    with torch.no_grad():
        ret = model.forward(batch['rays_o'], batch['rays_d'], batch['target_rgb'], batch['target_d'])
    assert sorted(ret) == sorted(['rgb', 'depth', 'disp_map', 'acc_map', 'depth_var', 'z_vals', 'raw'])
    for k in ('rays_o', 'rays_d', 'target_rgb', 'target_d'):
        out['eval_' + k] = batch[k].numpy()
    for k, v in ret.items():
        out['eval_ret_' + k] = v.numpy()

    # pose parameterisation
    aa2m, at2T = reference_pose_functions()
    rot = torch.randn(9, 3, generator=gen) * torch.tensor([0.02, 0.5, 1.5]).repeat(3)[:9, None]
    trans = torch.randn(9, 3, generator=gen)
    out.update(pose_rot=rot.numpy(), pose_trans=trans.numpy(), pose_R=aa2m(rot).numpy(), pose_T=at2T(rot, trans).numpy())

    path = os.path.join(ROOT, 'tests', 'golden', 'small_inference.npz')
    np.savez_compressed(path, **out)
    print('small_inference -> %s (%d KiB)' % (path, os.path.getsize(path) // 1024))


if __name__ == '__main__':
    main()
