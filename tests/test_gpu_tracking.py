"""Tracking on a frozen map (BASELINE.json: "tracking ATE within 5 % on a fixed synthetic trajectory").

A small map is first fitted on the GPU to an analytic box scene, then a short synthetic trajectory is tracked
twice from the same perturbed initial poses and with the same sample noise: by the product (fused kernels,
ray-gradient-only backward) and by the CPU oracle (torch autograd).  The two ATEs must agree within 5 %."""
import math

import pytest
import torch

from mneslam_b200 import configs, pose, synthetic
from mneslam_b200.mapping import FusedMapper
from mneslam_b200.tracking import PoseTracker, ate_rmse, matrix_to_quaternion
from oracle import joint_encoding as J
from tests.gpu_util import build_model, flat_planes

pytestmark = pytest.mark.gpu


def _camera_dirs(cfg, n, gen):
    H, W, fx, fy, cx, cy = 48, 64, 40.0, 40.0, 31.5, 23.5
    pix = torch.randperm(H * W, generator=gen)[:n]
    j, i = (pix // W).float(), (pix % W).float()
    return torch.stack([(i - cx) / fx, -(j - cy) / fy, -torch.ones_like(i)], -1)


def _pose(a):
    ang = 0.25 * math.sin(2.0 * a)
    R = pose.axis_angle_to_matrix(torch.tensor([[0.05, 1.0, 0.1]]) / 1.0099 * (ang + 1e-3))[0]
    c2w = torch.eye(4)
    c2w[:3, :3] = R
    c2w[:3, 3] = torch.tensor([-0.25 + 0.5 * a, 0.05 + 0.1 * math.sin(3 * a), 0.05 * math.cos(2 * a)])
    return c2w


@pytest.mark.parametrize('rot_rep', ['quat', 'axis_angle'])
def test_tracking_ate_matches_oracle(rot_rep):
    cfg = configs.small_test_config(n_samples_d=13, n_range_d=7)
    cfg['training']['rot_rep'] = rot_rep
    cfg['training']['range_d'] = 0.15
    bound = J.enlarged_bound(cfg)
    bb = torch.tensor(cfg['mapping']['bound'], dtype=torch.float64)
    gen = torch.Generator().manual_seed(5)
    model = build_model(cfg, bb, J.init_planes(cfg, bound, seed=4, std=0.01), J.init_weights(cfg, seed=4), as_params=False)

    # ---- fit the map on a few keyframes of the analytic scene
    mapper = FusedMapper(model)
    for it in range(400):
        c2w = _pose((it % 9) / 8.0)
        dirs = _camera_dirs(cfg, 512, gen)
        rgb, d = synthetic.render_box_scene(cfg, c2w, dirs)
        ro, rd = pose.rays_from_c2w(c2w, dirs)
        mapper.step(ro.contiguous().cuda(), rd.contiguous().cuda(), rgb.cuda(), d.cuda())
    torch.cuda.synchronize()
    planes_cpu = tuple([p.detach().cpu().contiguous() for p in grp] for grp in model.all_planes)
    weights_cpu = {k: w.detach().cpu() for k, w in zip(('sdf0', 'sdf1', 'col0', 'col1'), model.decoder.weight_list())}

    # ---- track a short trajectory, both implementations
    n_frames, n_iters, n_rays = 6, 20, 256
    S = synthetic.n_samples_total(cfg)
    gt, est_gpu, est_cpu = [], [], []
    t, m = cfg['training'], cfg['mapping']
    for f in range(n_frames):
        c2w_gt = _pose(0.1 + 0.8 * f / (n_frames - 1))
        init = c2w_gt.clone()
        init[:3, 3] += torch.tensor([0.03, -0.02, 0.025])
        init[:3, :3] = pose.axis_angle_to_matrix(torch.tensor([[0.02, -0.015, 0.01]]))[0] @ c2w_gt[:3, :3]
        dirs = _camera_dirs(cfg, n_rays, gen)
        rgb, d = synthetic.render_box_scene(cfg, c2w_gt, dirs)
        noises = [synthetic.make_noise(n_rays, S, seed=100 * f + i) for i in range(n_iters)]

        tracker = PoseTracker(model, init.cuda(), rot_rep=rot_rep, lr_rot=2e-3, lr_trans=2e-3)
        for i in range(n_iters):
            tracker.step(dirs.cuda(), rgb.cuda(), d.cuda(), noise=noises[i].cuda())
        est_gpu.append(tracker.trans.detach().cpu()[0])
        for p in flat_planes(model) + list(model.decoder.parameters()):
            assert p.grad is None or float(p.grad.abs().max()) == 0.0, 'the map must stay frozen while tracking'

        # oracle: same parameterisation, torch autograd on the CPU
        from mneslam_b200.tracking import matrix_to_axis_angle
        rot0 = matrix_to_quaternion(init[:3, :3]) if rot_rep == 'quat' else matrix_to_axis_angle(init[:3, :3])
        rot = rot0[None].clone().requires_grad_(True)
        trans = init[None, :3, 3].clone().requires_grad_(True)
        opt = torch.optim.Adam([{'params': [rot], 'lr': 2e-3}, {'params': [trans], 'lr': 2e-3}])
        for i in range(n_iters):
            opt.zero_grad()
            Rm = J.quat_to_matrix(rot)[0] if rot_rep == 'quat' else J.axis_angle_to_matrix(rot)[0]
            ro, rd = J.rays_from_pose(Rm, trans[0], dirs)
            ret, _ = J.forward(cfg, planes_cpu, weights_cpu, bound, bb, ro, rd, rgb, d, noises[i])
            J.total_loss(cfg, ret).backward()
            opt.step()
        est_cpu.append(trans.detach()[0])
        gt.append(c2w_gt[:3, 3])

    gt, est_gpu, est_cpu = torch.stack(gt), torch.stack(est_gpu), torch.stack(est_cpu)
    init_err = float(torch.linalg.norm(torch.tensor([0.03, -0.02, 0.025])))
    ate_gpu, ate_cpu = ate_rmse(est_gpu, gt), ate_rmse(est_cpu, gt)
    raw_gpu = float((est_gpu - gt).norm(dim=1).mean())
    assert raw_gpu < init_err, 'tracking must move the poses towards the ground truth (%g vs %g)' % (raw_gpu, init_err)
    assert abs(ate_gpu - ate_cpu) <= 0.05 * ate_cpu + 1e-5, (ate_gpu, ate_cpu)


def test_graphed_tracker_matches_eager():
    """With the stratified noise switched off the iteration is deterministic up to atomic ordering, so ten CUDA-graph
    replays must land on the same pose as ten eager steps; the map stays untouched."""
    from mneslam_b200.tracking import GraphedPoseTracker
    cfg = configs.small_test_config(n_samples_d=13, n_range_d=7)
    bound = J.enlarged_bound(cfg)
    bb = torch.tensor(cfg['mapping']['bound'], dtype=torch.float64)
    gen = torch.Generator().manual_seed(5)
    model = build_model(cfg, bb, J.init_planes(cfg, bound, seed=4, std=0.01), J.init_weights(cfg, seed=4), as_params=False)
    mapper = FusedMapper(model)
    for it in range(200):
        c2w = _pose((it % 9) / 8.0)
        dirs = _camera_dirs(cfg, 512, gen)
        rgb, d = synthetic.render_box_scene(cfg, c2w, dirs)
        ro, rd = pose.rays_from_c2w(c2w, dirs)
        mapper.step(ro.contiguous().cuda(), rd.contiguous().cuda(), rgb.cuda(), d.cuda())
    cfg['training']['perturb'] = 0                      # deterministic sampling from here on
    before = [p.clone() for p in flat_planes(model)]
    c2w_gt = _pose(0.4)
    init = c2w_gt.clone()
    init[:3, 3] += torch.tensor([0.03, -0.02, 0.025])
    dirs = _camera_dirs(cfg, 256, gen)
    rgb, d = synthetic.render_box_scene(cfg, c2w_gt, dirs)
    eager = PoseTracker(model, init.cuda(), rot_rep='quat', lr_rot=2e-3, lr_trans=2e-3)
    for _ in range(13):
        eager.step(dirs.cuda(), rgb.cuda(), d.cuda())
    graphed = PoseTracker(model, init.cuda(), rot_rep='quat', lr_rot=2e-3, lr_trans=2e-3)
    g = GraphedPoseTracker(graphed, dirs.cuda(), rgb.cuda(), d.cuda())      # three warm-up steps inside
    g.run(10)
    torch.cuda.synchronize()
    assert torch.allclose(graphed.trans, eager.trans, atol=2e-4), (graphed.trans, eager.trans)
    assert torch.allclose(graphed.rot, eager.rot, atol=2e-4), (graphed.rot, eager.rot)
    assert float((graphed.trans.detach().cpu()[0] - init[:3, 3]).norm()) > 1e-3, 'the pose must have moved'
    for a, b in zip(before, flat_planes(model)):
        assert torch.equal(a, b)


@pytest.mark.parametrize('rot_rep', ['quat', 'axis_angle'])
def test_fused_tracker_matches_autograd_tracker(rot_rep):
    """FusedTracker (mne_pose_rays + ray-gradient backward + mne_pose_step: dual-number chain rule and Adam in one kernel)
    against PoseTracker (torch autograd over the same fused render + torch.optim.Adam): same pose gradient on the first
    iteration, same pose after twelve; also as CUDA-graph replays."""
    from mneslam_b200.tracking import FusedTracker, GraphedPoseTracker
    cfg = configs.small_test_config(n_samples_d=13, n_range_d=7)
    bound = J.enlarged_bound(cfg)
    bb = torch.tensor(cfg['mapping']['bound'], dtype=torch.float64)
    gen = torch.Generator().manual_seed(5)
    model = build_model(cfg, bb, J.init_planes(cfg, bound, seed=4, std=0.3), J.init_weights(cfg, seed=4), as_params=False)
    cfg['training']['perturb'] = 0
    c2w_gt = _pose(0.4)
    init = c2w_gt.clone()
    init[:3, 3] += torch.tensor([0.03, -0.02, 0.025])
    dirs = _camera_dirs(cfg, 300, gen).cuda()
    rgb, d = synthetic.render_box_scene(cfg, c2w_gt, dirs.cpu())
    rgb, d = rgb.cuda(), d.cuda()
    ref = PoseTracker(model, init.cuda(), rot_rep=rot_rep, lr_rot=2e-3, lr_trans=2e-3)
    fused = FusedTracker(model, init.cuda(), rot_rep=rot_rep, lr_rot=2e-3, lr_trans=2e-3)
    assert torch.allclose(fused.c2w(), ref.c2w().detach(), atol=1e-6)
    l_ref = ref.step(dirs, rgb, d)
    l_fused = fused.step(dirs, rgb, d)
    torch.cuda.synchronize()
    assert l_fused.item() == pytest.approx(l_ref.item(), rel=1e-4)
    # PoseTracker leaves the gradients of its last backward on the parameters
    g_ref = torch.cat([ref.rot.grad.reshape(-1), ref.trans.grad.reshape(-1)])
    assert float((fused.pose_grad - g_ref).norm() / g_ref.norm()) < 2e-3
    for _ in range(11):
        ref.step(dirs, rgb, d)
        fused.step(dirs, rgb, d)
    torch.cuda.synchronize()
    assert torch.allclose(fused.trans, ref.trans.detach()[0], atol=2e-4), (fused.trans, ref.trans)
    assert torch.allclose(fused.rot, ref.rot.detach()[0], atol=2e-4), (fused.rot, ref.rot)
    assert float((fused.trans.cpu() - init[:3, 3]).norm()) > 1e-3, 'the pose must have moved'
    # graph replay of the fused iteration
    again = FusedTracker(model, init.cuda(), rot_rep=rot_rep, lr_rot=2e-3, lr_trans=2e-3)
    g = GraphedPoseTracker(again, dirs, rgb, d)          # three warm-up steps inside
    g.run(9)
    torch.cuda.synchronize()
    assert torch.allclose(again.trans, fused.trans, atol=2e-4) and torch.allclose(again.rot, fused.rot, atol=2e-4)


def test_fused_loop_alignment_matches_the_reference_loop():
    """FusedTracker.step_align against the loop of mp_slam/mapper.py:388-410 written with the drop-in class and torch:
    render_rays(target_d=None) -> rgb_weight * mse(rgb) + depth_weight * mse(depth) against a teacher rendering ->
    backward -> Adam on (axis-angle, translation)."""
    import torch.nn.functional as F
    from mneslam_b200.tracking import FusedTracker, frozen_map, matrix_to_axis_angle
    cfg = configs.small_test_config(n_samples_d=13, n_range_d=7, n_samples=24)
    cfg['training']['perturb'] = 0
    bound = J.enlarged_bound(cfg)
    bb = torch.tensor(cfg['mapping']['bound'], dtype=torch.float64)
    gen = torch.Generator().manual_seed(6)
    model = build_model(cfg, bb, J.init_planes(cfg, bound, seed=4, std=0.3), J.init_weights(cfg, seed=4), as_params=False)
    model.train()
    base = _pose(0.4)
    init = base.clone()
    init[:3, 3] += torch.tensor([0.02, -0.015, 0.01])
    dirs = _camera_dirs(cfg, 300, gen).cuda()
    with torch.no_grad():
        ro, rd = pose.rays_from_c2w(base.cuda(), dirs)
        teacher = model.render_rays(ro.contiguous(), rd.contiguous(), target_d=None)
        t_rgb, t_depth = teacher['rgb'].detach(), teacher['depth'].detach()
    # the reference loop (get_pose_param_optim: axis-angle for Replica-style configs)
    rot = torch.nn.Parameter(matrix_to_axis_angle(init[:3, :3].cuda())[None].clone())
    trans = torch.nn.Parameter(init[None, :3, 3].cuda().clone())
    opt = torch.optim.Adam([{'params': [rot], 'lr': 2e-3}, {'params': [trans], 'lr': 2e-3}])
    fused = FusedTracker(model, init.cuda(), rot_rep='axis_angle', lr_rot=2e-3, lr_trans=2e-3)
    w_c, w_d = cfg['training']['rgb_weight'], cfg['training']['depth_weight']
    for it in range(8):
        opt.zero_grad()
        c2w = pose.transform_matrix(rot, trans, 'axis_angle')[0]
        with frozen_map(model):
            ro, rd = pose.rays_from_c2w(c2w, dirs)
            r = model.render_rays(ro.contiguous(), rd.contiguous(), target_d=None)
            loss = w_c * F.mse_loss(r['rgb'], t_rgb) + w_d * F.mse_loss(r['depth'], t_depth)
            loss.backward()
        opt.step()
        l_fused = fused.step_align(dirs, t_rgb, t_depth)
        assert l_fused.item() == pytest.approx(loss.item(), rel=2e-3, abs=1e-9)
    torch.cuda.synchronize()
    assert torch.allclose(fused.trans, trans.detach()[0], atol=2e-4), (fused.trans, trans)
    assert torch.allclose(fused.rot, rot.detach()[0], atol=2e-4), (fused.rot, rot)
    assert float((fused.trans.cpu() - init[:3, 3]).norm()) > 1e-3
