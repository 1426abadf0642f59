#!/usr/bin/env python
"""Benchmark of the JointEncoding mapping iteration (BASELINE.json metric: sampled points/s, fwd+bwd map step).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--no-graph]

A "step" is ONE mapping iteration of the reference's hot loop (mp_slam/mapper.py:134-162) on the Replica
room0 configuration (BASELINE.json configs[1]): a 2048-ray batch drawn from a synthetic 20-keyframe ray
database, 43 samples per ray (88 064 points), forward + losses + backward (plane, decoder gradients) + Adam.
`value` = points / s of the whole job with all inputs resident in HBM (the step is captured in a CUDA graph);
`e2e` = the same iteration through the library's iteration driver (FusedMapper.step over the C ABI) with the ray batch
coming from pinned HOST memory and the loss read back to the host every step, no CUDA graph; `e2e_dropin` = the
drop-in JointEncoding class driven exactly like the reference's call sites (forward / loss.backward() /
torch.optim.Adam) on the same host buffers.  Multi-GPU (torchrun): every rank renders its own 2048-ray shard of a
global batch, loss accumulators and gradients are all-reduced over NCCL (weak scaling).
"""
import argparse
import json
import os
import subprocess
import sys
import tempfile
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

import torch  # noqa: E402

N_RAYS = 2048
N_KF = 20
MIN_PIXELS_CUR = {'room0': 100, 'scannet_hash': 20}      # mapping.min_pixels_cur of the reference's replica.yaml / scannet.yaml
ALG_BYTES_PER_POINT_FWD = 3072      # 3 planes x 2 levels x 4 texels x 32 ch x 4 B gathered      
ALG_BYTES_PER_POINT_BWD = 3072      # the same texels receive one gradient contribution each
METRIC = 'sampled points/s fwd+bwd (map step)'
UNIT = 'points/s'


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument('--gpus', type=int, default=1)
    ap.add_argument('--steps', type=int, default=50)
    ap.add_argument('--warmup', type=int, default=5)
    ap.add_argument('--impl', default='ours', choices=['ours', 'reference'])
    ap.add_argument('--no-graph', action='store_true')
    ap.add_argument('--no-pipeline', action='store_true', help='do not draw / sample the next batch under the optimiser of the current iteration')
    ap.add_argument('--no-cpu-baseline', action='store_true')
    ap.add_argument('--hash-fp16', action='store_true', help='scannet_hash workload: read the table through its fp16 copy (grid.hash_fp16)')
    ap.add_argument('--no-extras', action='store_true', help='skip the self_check / configs3 / gather_peak / steady-state records (profiling runs)')
    ap.add_argument('--rays', type=int, default=N_RAYS, help="keyframe rays per GPU and iteration (mapping.sample); the reference's batch adds "
                    "max(sample // n_keyframes, min_pixels_cur) rays of the current frame (mp_slam/mapper.py:137-139)")
    ap.add_argument('--no-current-frame', action='store_true', help='keyframe rays only (the round-1 batch)')
    ap.add_argument('--workload', default='room0', choices=['room0', 'scannet_hash'],
                    help='room0: BASELINE.json configs[1] (the default, the judged line); scannet_hash: configs[3] -- ScanNet scene0000 bound '
                         'and sampling (117 samples per ray) on the 16-level hash grid; use with --rays = 65536 / GPUs and --value-only')
    ap.add_argument('--no-mark-dirty', action='store_true', help='multi-GPU: derive the flags of the non-zero gradient texels from a pass over the gradient buffer instead of setting them in the backward kernel')
    ap.add_argument('--spare-sms', type=int, default=8, help='SMs the optimiser kernels leave to the pipelined side stream')
    ap.add_argument('--value-only', action='store_true', help='print the graph-replayed step time only (side experiments)')
    args = ap.parse_args()
    # the batch of mapping_optimize: `sample` rays over all keyframes + the current frame's share
    args.sample = args.rays
    args.cur = 0 if args.no_current_frame else max(args.rays // N_KF, MIN_PIXELS_CUR[args.workload])
    args.rays = args.sample + args.cur
    return args


def workload_config(cfg, n_rays, world):
    s = cfg['training']['n_samples_d'] + cfg['training']['n_range_d']
    return {'workload': 'Replica room0 mapping iteration: %d-ray batch/GPU (mapping.sample rays drawn without replacement from a synthetic '
                        '%d-keyframe DB, 1200x680, + max(sample // n_kf, min_pixels_cur) pixels of the current frame, as mapping_optimize '
                        'builds it), %d samples/ray, tri-plane 2x3 planes (129.4 MB fp32) + OneBlob + SDF/colour MLPs, '
                        'fwd + losses + bwd + Adam' % (n_rays, N_KF, s),
            'rays_per_gpu': n_rays, 'samples_per_ray': s, 'points_per_step': n_rays * s * world,
            'parallelism': 'ray-sharded data parallel x%d' % world if world > 1 else 'single GPU',
            'l2': 'inputs larger than L2: planes + gradients + Adam moments = 517 MB touched per step (L2 = 126 MB)'}


# ----------------------------------------------------------------------------------------------------
# CPU baseline: the oracle port of the reference's path (torch CPU ops incl. F.grid_sample, as the reference)
# ----------------------------------------------------------------------------------------------------
class CpuMapper:
    def __init__(self, cfg, seed=0, device='cpu'):
        from oracle import joint_encoding as J
        self.J, self.cfg = J, cfg
        self.bound = J.enlarged_bound(cfg)
        self.bb = torch.tensor(cfg['mapping']['bound'], dtype=torch.float64)
        planes = J.init_planes(cfg, self.bound, seed=seed, std=0.01)
        self.bound, self.bb = self.bound.to(device), self.bb.to(device)
        self.planes = tuple([torch.nn.Parameter(p.to(device)) for p in grp] for grp in planes)
        self.weights = {k: torch.nn.Parameter(v.to(device)) for k, v in J.init_weights(cfg, seed=seed).items()}
        mp = cfg['mapping']
        self.opt = torch.optim.Adam([
            {'params': list(self.weights.values()), 'weight_decay': 1e-6, 'lr': mp['lr_decoder']},
            {'params': [p for g in self.planes for p in g], 'eps': 1e-15, 'lr': mp['lr_embed']}], betas=(0.9, 0.99))

    def step(self, batch, noise):
        J = self.J
        ret, _ = J.forward(self.cfg, self.planes, self.weights, self.bound, self.bb, batch['rays_o'], batch['rays_d'],
                           batch['target_rgb'], batch['target_d'], noise, fast=True)
        loss = J.total_loss(self.cfg, ret)
        loss.backward()
        self.opt.step()
        self.opt.zero_grad()
        return float(loss.detach())


REF_STAGED = os.path.join(ROOT, 'oracle', '_ref')


class RefCpuMapper:
    """The reference's OWN modules (model/scene_rep.py::JointEncoding + decoder / encodings / utils, staged into the
    git-ignored oracle/_ref/ by __graft_entry__.build()) driven as its mapper drives them: forward, get_loss_from_ret
    (mneslam_mp.py:350-372), backward, the Adam of create_optimizer (mneslam_mp.py:431-469).  tiny-cuda-nn is absent from
    this image, so its one live object -- the OneBlob encoding -- is the oracle's restatement (oracle/tcnn_stub.py)."""

    def __init__(self, cfg, seed=0):
        import numpy as np
        from oracle import joint_encoding as J, tcnn_stub
        tcnn_stub.install()
        if REF_STAGED not in sys.path:
            sys.path.insert(0, REF_STAGED)
        from model.scene_rep import JointEncoding as RefJointEncoding          # the reference's class
        self.J, self.cfg = J, cfg
        torch.manual_seed(seed)
        import contextlib
        with contextlib.redirect_stdout(sys.stderr):       # the reference prints while it builds its encoders; stdout is the JSON line's
            self.model = RefJointEncoding(cfg, torch.from_numpy(np.array(cfg['mapping']['bound'])))
        self.model.train()
        planes = []
        for grp in self.model.all_planes:                                      # mneslam_mp.py:451-457
            for i, p in enumerate(grp):
                grp[i] = torch.nn.Parameter(p.cpu())    # (the reference puts its planes on the GPU when one is visible; this arm is the CPU path)
                planes.append(grp[i])
        mp = cfg['mapping']
        self.opt = torch.optim.Adam([{'params': self.model.decoder.parameters(), 'weight_decay': 1e-6, 'lr': mp['lr_decoder']},
                                     {'params': planes, 'eps': 1e-15, 'lr': mp['lr_embed']}], betas=(0.9, 0.99))

    def step(self, batch, noise):
        ret = self.model.forward(batch['rays_o'], batch['rays_d'], batch['target_rgb'], batch['target_d'])
        loss = self.J.total_loss(self.cfg, ret)
        loss.backward()
        self.opt.step()
        self.opt.zero_grad()
        return float(loss.detach())


def reference_staged():
    return os.path.exists(os.path.join(REF_STAGED, 'model', 'scene_rep.py'))


def time_cpu(cfg, n_rays, steps, warmup, threads=None):
    from mneslam_b200 import synthetic
    # torch.distributed.run exports OMP_NUM_THREADS=1: set the thread count explicitly, or every N > 1 launch would time a
    # single-threaded reference
    torch.set_num_threads(threads or os.cpu_count() or 1)
    s = synthetic.n_samples_total(cfg)
    m = RefCpuMapper(cfg) if reference_staged() else CpuMapper(cfg)
    times = []
    for i in range(warmup + steps):
        batch = synthetic.make_ray_batch(cfg, n_rays, seed=i)
        noise = synthetic.make_noise(n_rays, s, seed=i)
        t0 = time.perf_counter()
        m.step(batch, noise)
        dt = time.perf_counter() - t0
        if i >= warmup:
            times.append(dt)
    times.sort()
    med = times[len(times) // 2]
    return n_rays * s / med, med, sum(times)


def torch_gpu_record(cfg, n_rays, dev, steps=10, warmup=3):
    """The same oracle port with every tensor on the GPU: torch's own CUDA kernels (grid_sample, matmul, multi-tensor
    Adam) -- the operator mix of the reference's GPU path, OneBlob restated in torch instead of tiny-cuda-nn.  Reported
    beside cpu_baseline for orientation; device-resident inputs, CUDA events."""
    from mneslam_b200 import synthetic
    s = synthetic.n_samples_total(cfg)
    m = CpuMapper(cfg, device=dev)
    batches = [{k: v.to(dev) for k, v in synthetic.make_ray_batch(cfg, n_rays, seed=i).items()} for i in range(4)]
    noises = [synthetic.make_noise(n_rays, s, seed=i).to(dev) for i in range(4)]
    for i in range(warmup):
        m.step(batches[i % 4], noises[i % 4])
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for i in range(steps):
        m.step(batches[i % 4], noises[i % 4])
    b.record()
    torch.cuda.synchronize()
    ms = a.elapsed_time(b) / steps
    return {'value': n_rays * s / (ms * 1e-3), 'unit': UNIT, 'ms_per_step': ms, 'steps': steps,
            'path': 'oracle port on the GPU: torch CUDA ops (F.grid_sample, matmul, torch.optim.Adam), fp32, eager; includes one '
                    'loss.item() per step (the read-back of the e2e contract; the reference loop itself never reads the loss)'}


def cpu_kind():
    return 'reference' if reference_staged() else 'port'


def cpu_what():
    if reference_staged():
        return ("the reference's own modules (model/scene_rep.py JointEncoding.forward + get_loss_from_ret + backward + the Adam of "
                'create_optimizer), torch CPU, OneBlob restated (tiny-cuda-nn absent)')
    return "oracle port of the reference's torch-CPU path (F.grid_sample + matmul + torch Adam)"


def cpu_baseline_record(cfg, n_rays, steps=24, warmup=2):
    pts_s, med, total = time_cpu(cfg, n_rays, steps, warmup)
    threads = torch.get_num_threads()
    # the reference pins OMP_NUM_THREADS=3 (mneslam_mp.py:2): its own setting, reported beside the all-cores figure
    pts_s3, med3, _ = time_cpu(cfg, n_rays, 8, 1, threads=3)
    torch.set_num_threads(threads)
    return {'value': pts_s, 'unit': UNIT, 'cores': threads, 'kind': cpu_kind(),
            'sample': '%d mapping iterations (median) of the same %d-ray x %d-sample workload, %s, %d host threads of %d cores'
                      % (steps, n_rays, cfg['training']['n_samples_d'] + cfg['training']['n_range_d'], cpu_what(),
                         threads, os.cpu_count()),
            'ms_per_step': med * 1e3,
            'value_at_reference_threads': pts_s3, 'reference_threads': 3, 'ms_per_step_at_reference_threads': med3 * 1e3}


def run_reference(args):
    """--impl reference: the reference's CPU implementation of the path on the host cores, rank 0 only: its own modules
    where __graft_entry__.build() staged them (oracle/_ref/, git-ignored, travels with the snapshot), else the oracle port."""
    if int(os.environ.get('RANK', '0')) != 0:
        return
    from mneslam_b200 import configs
    cfg = configs.replica_room0()
    # the driver's own step / warm-up counts (one CPU iteration is ~0.3 s: K = 50, W = 5 is 15 s); bounded only against absurd values
    steps, warmup = max(1, min(args.steps, 200)), max(0, min(args.warmup, 50))
    pts_s, med, _ = time_cpu(cfg, args.rays, steps, warmup)
    sample = ('%d mapping iterations (median) after %d warm-up, %d rays x 43 samples, %s, %d host threads of %d cores'
              % (steps, warmup, args.rays, cpu_what(), torch.get_num_threads(), os.cpu_count()))
    line = {'impl': 'reference', 'metric': METRIC, 'value': pts_s, 'unit': UNIT, 'n_gpus': args.gpus, 'steps': steps,
            'warmup': warmup, 'ms_per_step': med * 1e3, 'higher_is_better': True, 'scaling': 'weak',
            'vs_baseline': None, 'dtype': 'f32', 'data': 'synthetic',
            'config': workload_config(cfg, args.rays, 1),
            'cpu_baseline': {'value': pts_s, 'unit': UNIT, 'cores': torch.get_num_threads(), 'kind': cpu_kind(),
                             'sample': sample},
            'e2e': {'value': pts_s, 'unit': UNIT, 'h2d_bytes_per_step': 0, 'd2h_bytes_per_step': 0},
            'gpu_launches': 0}
    print(json.dumps(line), flush=True)


# ----------------------------------------------------------------------------------------------------
# clocks
# ----------------------------------------------------------------------------------------------------
class ClockSampler:
    Q = ('clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,'
         'clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,'
         'clocks_event_reasons.sw_power_cap')

    def __init__(self, index):
        self.path = tempfile.mktemp(suffix='.csv')
        self.proc = None
        try:
            self.proc = subprocess.Popen(['nvidia-smi', '-i', str(index), '--query-gpu=' + self.Q,
                                          '--format=csv,noheader,nounits', '-lms', '100'],
                                         stdout=open(self.path, 'w'), stderr=subprocess.DEVNULL)
        except OSError:
            pass

    def stop(self):
        if self.proc is None:
            return {'sm_mhz': None, 'sm_max_mhz': None, 'reasons': ['nvidia-smi unavailable']}
        time.sleep(0.15)
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except Exception:
            self.proc.kill()
        sm, mx, reasons = [], None, set()
        names = ('hw_slowdown', 'hw_thermal_slowdown', 'sw_thermal_slowdown', 'sw_power_cap')
        for line in open(self.path):
            f = [x.strip() for x in line.split(',')]
            if len(f) < 7:
                continue
            try:
                sm.append(float(f[0])); mx = float(f[1])
            except ValueError:
                continue
            for n, v in zip(names, f[3:7]):
                if v.lower().startswith('active'):
                    reasons.add(n)
        os.unlink(self.path)
        sm.sort()
        return {'sm_mhz': sm[len(sm) // 2] if sm else None, 'sm_max_mhz': mx, 'samples': len(sm),
                'reasons': sorted(reasons)}


# ----------------------------------------------------------------------------------------------------
# our arm
# ----------------------------------------------------------------------------------------------------
def measured_peak():
    path = os.path.join(ROOT, 'MEASURED_PEAKS.json')
    if os.path.exists(path):
        return json.load(open(path))['hbm_gbs'], 'MEASURED_PEAKS.json hbm_gbs (measured)'
    return 6650.0, 'fallback 6.65 TB/s (B200_PROFILING.md)'


def ncu_traffic(kernel):
    """dram bytes per launch of a kernel from the committed ncu summary (profiles/traffic.json), if there is one."""
    path = os.path.join(ROOT, 'profiles', 'traffic.json')
    if os.path.exists(path):
        try:
            table = json.load(open(path)).get('dram_bytes_per_launch', {})
            for name, v in table.items():
                if name.startswith(kernel):
                    return v
        except Exception:
            return None
    return None



# ----------------------------------------------------------------------------------------------------
# records that travel in the same JSON line: BASELINE.json configs[3] and the multi-GPU self-check
# ----------------------------------------------------------------------------------------------------
CONFIGS3_TOTAL_RAYS = 65536
ALG_BYTES_PER_POINT_HASH = 2048     # 16 levels x 8 corners x 8 B gathered + the same entries receiving a gradient


def _max_over_ranks(ms, world, dev):
    if world > 1:
        import torch.distributed as dist
        t = torch.tensor([ms], device=dev, dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())
    return ms


def configs3_record(world, rank, dev, steps):
    """BASELINE.json configs[3]: 65 536 rays per iteration in TOTAL (strong scaling: 65 536 / N per GPU), ScanNet scene0000
    bound and sampling (96 + 21 samples per ray), 16-level hash grid (T = 2^19), gradients exchanged over NVLink; the same
    CUDA-graph path as the headline line.  Device time, max over ranks."""
    import torch.distributed as dist
    from mneslam_b200 import JointEncoding, configs, synthetic
    from mneslam_b200.mapping import DeviceKeyframeRays, FusedMapper
    cfg = configs.scannet_scene0000_onegrid()
    cfg['grid']['use_hash'] = True
    n = CONFIGS3_TOTAL_RAYS // world
    s = synthetic.n_samples_total(cfg)
    torch.manual_seed(0)
    model = JointEncoding(cfg, torch.tensor(cfg['mapping']['bound'], dtype=torch.float64)).to(dev)
    model.train()
    mapper = FusedMapper(model, world_size=world)
    rays_cpu, poses_cpu = synthetic.make_keyframe_db(cfg, n_kf=N_KF, seed=0)
    db = DeviceKeyframeRays(rays_cpu.to(dev), poses_cpu.to(dev))

    def one_step():
        return mapper.step(*db.sample(n))
    for _ in range(3):
        one_step()
    torch.cuda.synchronize()
    graph = None
    try:
        side = torch.cuda.Stream()
        side.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(side):
            one_step()
        torch.cuda.current_stream().wait_stream(side)
        torch.cuda.synchronize()
        graph = torch.cuda.CUDAGraph()
        with torch.cuda.graph(graph):
            one_step()
        graph.replay()
        torch.cuda.synchronize()
    except Exception as e:
        print('bench.py: configs3 graph capture failed (%r); eager' % (e,), file=sys.stderr)
        graph = None
        torch.cuda.synchronize()
    k = max(3, min(steps, 10))
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    e0.record()
    for _ in range(k):
        graph.replay() if graph is not None else one_step()
    e1.record()
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    ms = _max_over_ranks(e0.elapsed_time(e1) / k, world, dev)
    pts = n * s * world
    pts_s = pts / (ms * 1e-3)
    peak, _ = measured_peak()
    rec = {'workload': 'BASELINE.json configs[3]: %d rays per iteration in total (%d per GPU), ScanNet scene0000 bound, %d samples/ray, '
                       '16-level hash grid T=2^19 + OneBlob + MLPs, fwd + losses + bwd + Adam, gradient exchange over NVLink peer memory'
                       % (n * world, n, s),
           'scaling': 'strong', 'n_gpus': world, 'rays_total': n * world, 'rays_per_gpu': n, 'samples_per_ray': s,
           'ms_per_step': ms, 'value': pts_s, 'unit': UNIT, 'steps': k, 'cuda_graph': graph is not None,
           'algorithmic_bytes_per_point': ALG_BYTES_PER_POINT_HASH,
           'gather_roofline_frac': pts_s * ALG_BYTES_PER_POINT_HASH / 1e9 / (peak * world),
           'table_entries': int(model.hash_meta.n_entries)}
    del graph, mapper, model, db
    torch.cuda.empty_cache()
    return rec


def self_check_record(world, rank, dev, iters=4):
    """Do the ranks compute the same map, and is it the single-GPU map?  Room0 configuration, 256 rays per rank: a few
    sharded iterations (flat parameter buffers must be bit-identical across ranks afterwards), against one GPU (rank 0)
    taking the same iterations on the concatenated batch."""
    import torch.distributed as dist
    from mneslam_b200 import JointEncoding, configs, synthetic
    from mneslam_b200.mapping import FusedMapper
    cfg = configs.replica_room0()
    bb = torch.tensor(cfg['mapping']['bound'], dtype=torch.float64)
    s = synthetic.n_samples_total(cfg)
    n = 256

    def make(world_size):
        torch.manual_seed(7)
        m = JointEncoding(cfg, bb).to(dev)
        m.train()
        return m, FusedMapper(m, world_size=world_size)
    model, mapper = make(world)
    single = make(1) if (rank == 0 and world > 1) else None
    loss_err = 0.0
    for it in range(iters):
        shards = [{k: v.to(dev) for k, v in synthetic.make_ray_batch(cfg, n, seed=500 + it * 64 + r, zero_depth_frac=0.05).items()}
                  for r in (range(world) if rank == 0 else [rank])]
        noises = [synthetic.make_noise(n, s, seed=900 + it * 64 + r).to(dev) for r in (range(world) if rank == 0 else [rank])]
        mine, nz = (shards[rank], noises[rank]) if rank == 0 else (shards[0], noises[0])
        losses = mapper.step(mine['rays_o'], mine['rays_d'], mine['target_rgb'], mine['target_d'], noise=nz)
        if single is not None:
            cat = {k: torch.cat([b[k] for b in shards]) for k in ('rays_o', 'rays_d', 'target_rgb', 'target_d')}
            ref = single[1].step(cat['rays_o'], cat['rays_d'], cat['target_rgb'], cat['target_d'], noise=torch.cat(noises))
            a, b = losses[:7].double(), ref[:7].double()
            ok = torch.isfinite(b)
            loss_err = max(loss_err, float(((a[ok] - b[ok]).abs() / (b[ok].abs() + 1e-12)).max()))
    torch.cuda.synchronize()
    identical = True
    if world > 1:
        flat = mapper.flat_param.clone()
        dist.broadcast(flat, src=0)
        same = torch.tensor([1 if torch.equal(flat, mapper.flat_param) else 0], device=dev)
        dist.all_reduce(same, op=dist.ReduceOp.MIN)
        identical = bool(int(same.item()))
    err = None
    if single is not None:
        a = mapper.flat_param[:mapper.total].double()
        b = single[1].flat_param[:single[1].total].double()
        err = float((a - b).norm() / (b.norm() + 1e-30))
        moved = float((b - b.new_zeros(()).expand_as(b)).norm())      # sanity: the map is not all zeros
        err_ok = err < 2e-3 and moved > 0
    else:
        err_ok = True
    timeouts = mapper.barrier_timeouts() if hasattr(mapper, 'barrier_timeouts') else 0
    rec = {'ok': bool(identical and err_ok and loss_err < 1e-3 and timeouts == 0), 'ranks_identical': identical, 'barrier_timeouts': timeouts,
           'param_rel_err_vs_single_gpu': err, 'loss_rel_err_vs_single_gpu': loss_err if single is not None else None,
           'iterations': iters, 'rays_per_rank': n,
           'what': 'room0 map, %d sharded iterations on %d ranks; flat parameter buffers compared bit for bit across ranks '
                   '(broadcast + equal) and against one GPU taking the concatenated batches (relative L2 error < 2e-3: the '
                   'summation order of the float atomics differs)' % (iters, world)}
    if world > 1:
        ok_all = torch.tensor([1 if rec['ok'] else 0], device=dev)
        dist.broadcast(ok_all, src=0)
    del mapper, model, single
    torch.cuda.empty_cache()
    return rec


def gather_peak_record(model, mapper, cfg, db, n_rays, dev):
    """The gather rate the GPU sustains on the batch's OWN tap addresses with a kernel that only gathers (mne_bench_gather):
    the L2 / HBM gather ceiling of SURVEY.md 8(d).  Two states: straight after a pass over a buffer larger than L2 (what a
    mapping iteration's forward finds: the optimiser has just streamed 517 MB), and warm (the same addresses again)."""
    import ctypes as C
    from mneslam_b200 import _lib, ops, synthetic
    lib = _lib.lib()
    s = synthetic.n_samples_total(cfg)
    ro, rd, trgb, td = db.sample(n_rays)
    tables = ops.SampleTables.get(cfg, dev, True)
    rcfg = ops.make_render_cfg(cfg, norm_rays=n_rays)
    z = ops.sample_rays(ops.make_sample_cfg(tables, True), rcfg, n_rays, s, td.reshape(-1), torch.rand(n_rays, s, device=dev), ops.new_acc(dev), dev)
    pts = (ro[:, None, :] + rd[:, None, :] * z[..., None]).reshape(-1, 3)
    b = model.bound.to(dev)
    nrm = (pts - b[:, 0]) / (b[:, 1] - b[:, 0]) * 2 - 1
    planes = model._flat_planes()
    base = mapper.flat_param
    offs = []
    for l in range(2):                                   # the forward's order: level, plane (xy, xz, yz), corner
        for a in range(3):
            pl = planes[a * 2 + l]
            H, W = pl.shape[2], pl.shape[3]
            gx = nrm[:, 1] if a == 2 else nrm[:, 0]
            gy = nrm[:, 1] if a == 0 else nrm[:, 2]
            ux = ((gx + 1) * 0.5 * (W - 1)).clamp(0, W - 1)
            uy = ((gy + 1) * 0.5 * (H - 1)).clamp(0, H - 1)
            x0, y0 = ux.floor().long(), uy.floor().long()
            x1, y1 = (x0 + 1).clamp(max=W - 1), (y0 + 1).clamp(max=H - 1)
            p0 = (pl.data_ptr() - base.data_ptr()) // 4
            for yy, xx in ((y0, x0), (y0, x1), (y1, x0), (y1, x1)):
                offs.append(p0 + (yy * W + xx) * 32)
    offsets = torch.stack(offs, 1).reshape(-1).to(torch.int32).contiguous()          # [M * 24]
    sink = torch.zeros(4, device=dev)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
    st = ops._stream()

    def run():
        _lib.check(lib.mne_bench_gather(base.data_ptr(), offsets.data_ptr(), offsets.numel(), 0, sink.data_ptr(), st), 'mne_bench_gather')
    for _ in range(3):
        run()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    cold = []
    for _ in range(10):
        flush.zero_()
        e0.record(); run(); e1.record()
        torch.cuda.synchronize()
        cold.append(e0.elapsed_time(e1))
    e0.record()
    for _ in range(20):
        run()
    e1.record()
    torch.cuda.synchronize()
    warm_ms = e0.elapsed_time(e1) / 20
    cold.sort()
    nbytes = offsets.numel() * 128
    return {'pattern': 'the 24 texel fetches (128 B each) of every sample of a %d-ray room0 batch, addresses computed from the batch' % n_rays,
            'bytes': nbytes, 'us_after_l2_flush': cold[len(cold) // 2] * 1e3, 'us_warm': warm_ms * 1e3,
            'gather_gbs_after_l2_flush': nbytes / (cold[len(cold) // 2] * 1e-3) / 1e9, 'gather_gbs_warm': nbytes / (warm_ms * 1e-3) / 1e9,
            'kernel': 'mne_bench_gather mode 0: loads only, 12 x 128-bit loads in flight per lane, 8 CTAs x 256 threads per SM'}


def exchange_record(world, rank, dev):
    """Agent-to-agent map exchange (BASELINE.json configs[4]): every rank holds a room0 map of its own and
    hands it to the next rank round a ring (mneslam_b200.exchange.exchange_maps: NCCL send / receive straight between the
    parameter storages), against the reference's mechanism -- torch.save of latest_checkpoint.pt, torch.load by the peer
    (mneslam_mp.py:294-315, mp_slam/mapper.py:708-726) -- timed on rank 0 in a local temporary directory."""
    import torch.distributed as dist
    from mneslam_b200 import JointEncoding, configs, exchange
    cfg = configs.replica_room0()
    bb = torch.tensor(cfg['mapping']['bound'], dtype=torch.float64)
    torch.manual_seed(100 + rank)
    model = JointEncoding(cfg, bb).to(dev)
    torch.manual_seed(0)
    shared = JointEncoding(cfg, bb).to(dev)
    nxt, prv = (rank + 1) % world, (rank - 1) % world
    nbytes = exchange.exchange_maps(model, shared, nxt, recv_from=prv)
    torch.cuda.synchronize(); dist.barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    times = []
    for _ in range(7):
        dist.barrier(); torch.cuda.synchronize()
        e0.record()
        exchange.exchange_maps(model, shared, nxt, recv_from=prv)
        e1.record()
        torch.cuda.synchronize()
        times.append(e0.elapsed_time(e1))
    ms = _max_over_ranks(sorted(times)[len(times) // 2], world, dev)
    want = torch.stack([t.detach().double().sum() for t in model._flat_planes()])
    sums = [torch.zeros_like(want) for _ in range(world)]
    dist.all_gather(sums, want)
    got = torch.stack([t.detach().double().sum() for t in shared._flat_planes()])
    ok = torch.tensor([1 if torch.equal(got, sums[prv]) else 0], device=dev)
    dist.all_reduce(ok, op=dist.ReduceOp.MIN)
    # the same as peer stores straight into model_shared (PeerMapInbox: it lives in NVLink peer-mapped memory)
    push_ms = push_ok = None
    try:
        torch.manual_seed(0)
        shared2 = JointEncoding(cfg, bb).to(dev)
        inbox = exchange.PeerMapInbox(shared2)
        exchange.push_map(model, inbox, nxt)
        torch.cuda.synchronize(); dist.barrier()
        times = []
        for _ in range(7):
            dist.barrier(); torch.cuda.synchronize()
            e0.record()
            exchange.push_map(model, inbox, nxt)
            e1.record()
            torch.cuda.synchronize()
            times.append(e0.elapsed_time(e1))
        push_ms = _max_over_ranks(sorted(times)[len(times) // 2], world, dev)
        got2 = torch.stack([t.detach().double().sum() for t in shared2._flat_planes()])
        ok2 = torch.tensor([1 if torch.equal(got2, sums[prv]) else 0], device=dev)
        dist.all_reduce(ok2, op=dist.ReduceOp.MIN)
        push_ok = bool(int(ok2.item()))
        del inbox, shared2
    except Exception as e:
        print('bench.py: peer-store exchange unavailable (%r)' % (e,), file=sys.stderr)
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
    del model, shared
    torch.cuda.empty_cache()
    best = min(x for x in (ms, push_ms) if x is not None)
    return {'ms': best, 'bytes_per_direction': nbytes, 'gb_per_s_per_direction': nbytes / (best * 1e-3) / 1e9,
            'ms_nccl_send_recv': ms, 'ms_peer_stores': push_ms, 'received_map_identical': bool(int(ok.item())) and push_ok is not False,
            'pattern': 'ring over %d ranks: every rank sends its 129 MB room0 map to the next and receives the previous one into '
                       'model_shared; device time, median of 7, max over ranks' % world,
            'reference_file_round_trip_ms': file_ms,
            'reference_mechanism': 'torch.save(latest_checkpoint.pt) + torch.load by the peer, local temporary directory, rank 0'}


def inference_record(model, cfg, dev):
    """The forward-only consumers of the same kernels: render_img as save_imgs calls it twice per keyframe
    (mneslam_mp.py:498,516: 1200 x 680 rays, 256 uniform samples without depth, 32 + 11 with), and a 500 000-point
    query_sdf chunk as the mesher issues them (model/utils.py:64-153).  Wall clock around the public methods, device
    synchronised; points/s and the fraction of the 3 072 B/point gather roofline."""
    from mneslam_b200 import synthetic
    cam = cfg['cam']
    H, W = cam['H'] - 2 * cam['crop_edge'], cam['W'] - 2 * cam['crop_edge']
    gen = torch.Generator().manual_seed(5)
    c2w = synthetic.make_pose(cfg, gen)
    gt_depth = (torch.rand(H, W, generator=gen) * 3.0 + 0.5).to(dev)
    peak, _ = measured_peak()
    was_training = model.training
    model.eval()
    out = {}
    for name, depth, s in (('render_img_no_depth', None, cfg['training']['n_samples']),
                           ('render_img_with_depth', gt_depth, cfg['training']['n_samples_d'] + cfg['training']['n_range_d'])):
        model.render_img(c2w, dev, gt_depth=depth)
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        model.render_img(c2w, dev, gt_depth=depth)
        torch.cuda.synchronize()
        ms = (time.perf_counter() - t0) * 1e3
        pts = H * W * s
        out[name] = {'ms': ms, 'rays': H * W, 'samples_per_ray': s, 'points': pts, 'points_per_s': pts / (ms * 1e-3),
                     'gather_roofline_frac': pts / (ms * 1e-3) * ALG_BYTES_PER_POINT_FWD / 1e9 / peak}
    pts = (torch.rand(500000, 1, 3, device=dev) * 2 - 1) * 0.9
    b = torch.tensor(cfg['mapping']['bound'], dtype=torch.float32, device=dev)
    pts = b[:, 0] + (pts * 0.5 + 0.5) * (b[:, 1] - b[:, 0])
    with torch.no_grad():
        model.query_sdf(pts)
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        for _ in range(5):
            model.query_sdf(pts)
        torch.cuda.synchronize()
    ms = (time.perf_counter() - t0) * 1e3 / 5
    out['query_sdf_500k'] = {'ms': ms, 'points': 500000, 'points_per_s': 500000 / (ms * 1e-3),
                             'gather_roofline_frac': 500000 / (ms * 1e-3) * ALG_BYTES_PER_POINT_FWD / 1e9 / peak}
    out['note'] = ('gather_roofline_frac = points/s x 3072 algorithmic bytes per point over the HBM copy bandwidth of MEASURED_PEAKS.json; '
                   'it passes 1 where neighbouring samples of a ray share texels and the fetches are served by L1/L2 instead of HBM')
    model.train(was_training)
    return out


def run_ours(args):
    import torch.distributed as dist
    from mneslam_b200 import JointEncoding, configs, synthetic
    from mneslam_b200.mapping import DeviceKeyframeRays, FusedMapper, loss_weight_vector
    from mneslam_b200 import ops

    rank = int(os.environ.get('RANK', '0'))
    world = int(os.environ.get('WORLD_SIZE', '1'))
    local = int(os.environ.get('LOCAL_RANK', '0'))
    if not torch.cuda.is_available():
        raise SystemExit('bench.py: no CUDA device (the product path has no CPU fallback); use --impl reference for the CPU arm')
    torch.cuda.set_device(local)
    dev = torch.device('cuda', local)
    if world > 1:
        if os.environ.get('NCCL_DEBUG', 'VERSION').upper() == 'VERSION':
            os.environ['NCCL_DEBUG'] = 'WARN'        # keep NCCL's version banner out of the one-JSON-line stdout
        dist.init_process_group('nccl', device_id=dev)

    if args.workload == 'scannet_hash':
        cfg = configs.scannet_scene0000_onegrid()
        cfg['grid']['use_hash'] = True
        cfg['grid']['hash_fp16'] = bool(args.hash_fp16)
    else:
        cfg = configs.replica_room0()
    n_rays = args.rays
    s = synthetic.n_samples_total(cfg)
    pts_per_rank = n_rays * s

    torch.manual_seed(0)
    bb = torch.tensor(cfg['mapping']['bound'], dtype=torch.float64)
    model = JointEncoding(cfg, bb).to(dev)
    model.train()
    # the map is replicated: same seed on every rank gives identical planes and decoder
    mapper = FusedMapper(model, world_size=world, mark_dirty=False if args.no_mark_dirty else None)
    rays_cpu, poses_cpu = synthetic.make_keyframe_db(cfg, n_kf=N_KF, seed=0)
    current_cpu = synthetic.make_current_frame(cfg, poses_cpu[-1], seed=rank) if args.cur else None
    db = DeviceKeyframeRays(rays_cpu.to(dev), poses_cpu.to(dev), current=None if current_cpu is None else current_cpu.to(dev),
                            min_pixels_cur=cfg['mapping']['min_pixels_cur'])
    n_sample = args.sample
    assert db.batch_rays(n_sample) == n_rays, (db.batch_rays(n_sample), n_rays)

    class _Draw:
        """db.sample for the helpers below that think in whole batches."""
        def sample(self, n):
            assert n == n_rays
            return db.sample(n_sample)
    batch_db = _Draw()

    # draw_rays | sample | points_fwd | rays_tail | loss_finalize | points_bwd3 | decoder_wgrad | adam_tick, adam_segments x2
    launches_per_step = 10

    def one_step():
        ro, rd, trgb, td = db.sample(n_sample)
        return mapper.step(ro, rd, trgb, td)

    # ---- warm-up (also sets kernel attributes outside capture)
    for _ in range(max(args.warmup, 3)):
        losses = one_step()
    torch.cuda.synchronize()

    graph = None
    if not args.no_graph:
        # the whole iteration -- ray assembly, five kernels' worth of rendering, NCCL collectives, Adam -- in one graph
        try:
            side = torch.cuda.Stream()
            side.wait_stream(torch.cuda.current_stream())
            with torch.cuda.stream(side):
                one_step()
            torch.cuda.current_stream().wait_stream(side)
            torch.cuda.synchronize()
            graph = torch.cuda.CUDAGraph()
            with torch.cuda.graph(graph):
                static_losses = one_step()
            for _ in range(3):
                graph.replay()
            torch.cuda.synchronize()
        except Exception as e:                      # e.g. a collective that cannot be captured: time it eagerly
            print('bench.py: CUDA-graph capture failed (%r); timing eager launches' % (e,), file=sys.stderr)
            graph = None
            torch.cuda.synchronize()

    # the same iteration software-pipelined: batch t+1 is drawn, assembled and sampled on a second stream while the optimiser of
    # iteration t runs (mapping.PipelinedMapper); every iteration does the same work
    pipe = None
    if graph is not None and not args.no_pipeline:
        try:
            from mneslam_b200.mapping import PipelinedMapper
            pipe = PipelinedMapper(mapper, db, n_sample, spare_sms=args.spare_sms).capture()
            for _ in range(3):
                pipe.step()
            torch.cuda.synchronize()
            static_losses = pipe.losses
        except Exception as e:
            print('bench.py: pipelined graph unavailable (%r); plain graph' % (e,), file=sys.stderr)
            pipe = None
            torch.cuda.synchronize()

    def run_steps(k):
        if pipe is not None:
            for _ in range(k):
                pipe.step()
        elif graph is not None:
            for _ in range(k):
                graph.replay()
        else:
            for _ in range(k):
                one_step()

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    # ---- timed region: exactly K steps, device time via CUDA events, max over ranks
    sampler = ClockSampler(local) if rank == 0 else None
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier()
    e0.record()
    run_steps(args.steps)
    e1.record()
    barrier()
    ms_total = e0.elapsed_time(e1)
    clocks = sampler.stop() if sampler else None
    if world > 1:
        t = torch.tensor([ms_total], device=dev, dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms_total = float(t.item())
    ms_step = ms_total / args.steps
    value = pts_per_rank * world / (ms_step * 1e-3)
    if args.value_only:
        if rank == 0:
            print(json.dumps({'metric': METRIC, 'value': value, 'unit': UNIT, 'n_gpus': world, 'steps': args.steps,
                              'ms_per_step': ms_step, 'workload': args.workload, 'rays_per_gpu': n_rays, 'samples_per_ray': s,
                              'cuda_graph': graph is not None, 'pipelined': pipe is not None, 'clocks': clocks,
                              'touched_fraction': float(mapper.touched.float().mean().item()) if mapper.touched is not None else None}),
                  flush=True)
        if world > 1:
            torch.cuda.synchronize(); dist.barrier(); torch.cuda.synchronize()
            sys.stdout.flush(); os._exit(0)
        return

    # ---- per-kernel device times (events around each launch, same stream, un-graphed), K steps
    def timed(fn):
        # a short device-side sleep first, so that the host has enqueued the launches before the first event is reached
        # on the stream: the interval then holds kernel time, not launch gaps
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        torch.cuda._sleep(400000)
        a.record(); r = fn(); b.record()
        return r, (a, b)
    tables = ops.SampleTables.get(cfg, dev, True)
    scfg = ops.make_sample_cfg(tables, True)
    rcfg = ops.make_render_cfg(cfg, norm_rays=n_rays * world)
    ev = {'sample': [], 'fwd': [], 'fwd_single_launch': [], 'bwd': [], 'wgrad': [], 'adam': []}
    counters = torch.zeros(n_rays, dtype=torch.int32, device=dev)
    k_prof = min(args.steps, 20)
    for _ in range(k_prof):
        ro, rd, trgb, td = db.sample(n_sample)
        tdf = td.reshape(-1)
        noise = torch.rand(n_rays, s, device=dev)
        acc = ops.new_acc(dev)
        z, e = timed(lambda: ops.sample_rays(scfg, rcfg, n_rays, s, tdf, noise, acc, dev)); ev['sample'].append(e)
        if world > 1:
            dist.all_reduce(acc[:6])
        # descriptors built outside the timed intervals
        sc_f, sc_b, sc_w = mapper._scene(False), mapper._scene(True, weight_grads=False), mapper._scene(True, plane_grads=False)
        scr = mapper._scratch
        # fwd = the fused forward: encode, decode, compositing, loss sums, d loss / d raw + compaction (one launch)
        out, e = timed(lambda: ops.render_forward(sc_f, rcfg, ro, rd, z, trgb, tdf, acc, want_aux=False, act=mapper._act,
                                                  ray_counters=None, loss_grads=mapper.loss_grads, scratch=scr)); ev['fwd'].append(e)
        out, e = timed(lambda: ops.render_forward(sc_f, rcfg, ro, rd, z, trgb, tdf, acc.clone(), want_aux=False, act=mapper._act,
                                                  ray_counters=counters, loss_grads=mapper.loss_grads, scratch=scr)); ev['fwd_single_launch'].append(e)
        # bwd = the point backward (decoder backward + plane scatter); wgrad = the decoder weight gradients, full grid
        _, e = timed(lambda: ops.render_backward_points(sc_b, ro, rd, z, scr, act=mapper._act)); ev['bwd'].append(e)
        _, e = timed(lambda: ops.decoder_wgrad(sc_w, n_rays, s, scr, mapper._act)); ev['wgrad'].append(e)
        mapper._pending_wgrad = None
        nz_frac = float((mapper.flat_grad[:mapper.dense_from].view(-1, 4) != 0).any(1).float().mean().item())
        _, e = timed(mapper.optimizer_step); ev['adam'].append(e)
    torch.cuda.synchronize()
    us = {k: 1e3 * sum(a.elapsed_time(b) for a, b in v) / len(v) for k, v in ev.items()}
    # ---- the optimiser late in a run: the fraction of parameter groups that ever received a gradient keeps rising for a few
    #      hundred iterations (0.43 -> 0.6), and the Adam pass with it; 400 more iterations, then the same interval again
    touched_prof = float(mapper.touched.float().mean().item()) if mapper.touched is not None else 1.0
    steady = None
    if world == 1 and not args.no_extras:
        run_steps(400)
        ev2 = []
        for _ in range(10):
            mapper.forward_backward(*db.sample(n_sample), defer_wgrad=True)
            _, e = timed(mapper.optimizer_step); ev2.append(e)
        torch.cuda.synchronize()
        steady = {'iterations_before': args.steps + 400 + 3 * max(args.warmup, 3), 'adam_us': 1e3 * sum(a.elapsed_time(b) for a, b in ev2) / len(ev2),
                  'touched_fraction': float(mapper.touched.float().mean().item()) if mapper.touched is not None else 1.0}

    # ---- e2e: the drop-in class through the reference's call sites, host buffers in, loss out (every step)
    planes = model._flat_planes()
    for grp in model.all_planes:
        for i, p in enumerate(grp):
            grp[i] = torch.nn.Parameter(p)                                  # mneslam_mp.py:451-457
    mp = cfg['mapping']
    opt = torch.optim.Adam([{'params': model.decoder.parameters(), 'weight_decay': 1e-6, 'lr': mp['lr_decoder']},
                            {'params': model._flat_planes(), 'eps': 1e-15, 'lr': mp['lr_embed']}], betas=(0.9, 0.99))
    lw = loss_weight_vector(cfg)
    n_pool = 8
    host_rays = [torch.empty(n_rays, 7).pin_memory() for _ in range(n_pool)]
    host_ids = [torch.empty(n_rays, dtype=torch.int64).pin_memory() for _ in range(n_pool)]
    g_cpu = torch.Generator().manual_seed(99 + rank)
    flat_db = rays_cpu.reshape(-1, 7)
    for hr, hi in zip(host_rays, host_ids):
        idx = torch.randperm(flat_db.shape[0], generator=g_cpu)[:n_sample]
        hr[:n_sample].copy_(flat_db[idx]); hi[:n_sample].copy_(idx // rays_cpu.shape[1])
        if args.cur:
            hr[n_sample:].copy_(current_cpu[torch.randperm(current_cpu.shape[0], generator=g_cpu)[:args.cur]]); hi[n_sample:].fill_(-1)
    h2d_bytes = n_rays * 7 * 4 + n_rays * 8
    loss_host = torch.empty(1).pin_memory()

    def e2e_step(i):
        rays7 = host_rays[i % n_pool].to(dev, non_blocking=True)
        ids = host_ids[i % n_pool].to(dev, non_blocking=True)
        ro, rd, trgb, td = db.assemble(rays7, ids)
        ret = model.forward(ro, rd, trgb, td)
        loss = (lw[0] * ret['rgb_loss'] + lw[1] * ret['depth_loss'] + lw[4] * ret['e_fs_loss']
                + lw[5] * ret['e_center_loss'] + lw[6] * ret['e_tail_loss'])
        loss.backward()
        if world > 1:
            for p in model._flat_planes() + list(model.decoder.parameters()):
                dist.all_reduce(p.grad)
        opt.step()
        opt.zero_grad()
        loss_host.copy_(loss.detach().reshape(1), non_blocking=False)      # D2H read of the step's loss
        return float(loss_host[0])

    # the same call sites with the loss read back one iteration late (the reference loop itself never reads it: mapper.py:155-162):
    # the host does not wait for the iteration it has just queued
    lag_host = [torch.empty(1).pin_memory() for _ in range(2)]
    lag_done = [torch.cuda.Event() for _ in range(2)]
    lag_state = {'n': 0}

    def e2e_step_lagged(i):
        rays7 = host_rays[i % n_pool].to(dev, non_blocking=True)
        ids = host_ids[i % n_pool].to(dev, non_blocking=True)
        ro, rd, trgb, td = db.assemble(rays7, ids)
        ret = model.forward(ro, rd, trgb, td)
        loss = (lw[0] * ret['rgb_loss'] + lw[1] * ret['depth_loss'] + lw[4] * ret['e_fs_loss']
                + lw[5] * ret['e_center_loss'] + lw[6] * ret['e_tail_loss'])
        loss.backward()
        if world > 1:
            for p in model._flat_planes() + list(model.decoder.parameters()):
                dist.all_reduce(p.grad)
        opt.step()
        opt.zero_grad()
        k = lag_state['n'] % 2
        lag_state['n'] += 1
        lag_host[k].copy_(loss.detach().reshape(1), non_blocking=True)      # D2H read of this step's loss, collected by the next call
        lag_done[k].record()
        if lag_state['n'] == 1:
            return None
        lag_done[1 - k].synchronize()
        return float(lag_host[1 - k][0])

    def e2e_lagged_finish():
        if lag_state['n']:
            k = (lag_state['n'] - 1) % 2
            lag_done[k].synchronize()
            lag_state['n'] = 0
            return float(lag_host[k][0])

    def fused_e2e_step(i):
        # the repo's own iteration driver fed from HOST memory: H2D of the ray batch, one iteration, D2H of the loss
        rays7 = host_rays[i % n_pool].to(dev, non_blocking=True)
        ids = host_ids[i % n_pool].to(dev, non_blocking=True)
        ro, rd, trgb, td = db.assemble(rays7, ids)
        losses = mapper.step(ro, rd, trgb, td)
        loss_host.copy_(mapper.total_loss(losses).reshape(1), non_blocking=False)
        return float(loss_host[0])

    hosted = None

    def graphed_e2e_step(i):
        # the same iteration as ONE graph launch: host memcpy into the pinned staging buffer, graph (H2D, assembly, iteration,
        # D2H of the loss), stream sync, loss as a Python float
        return hosted.step(host_rays[i % n_pool], host_ids[i % n_pool])

    def time_e2e(fn, finish=None):
        for i in range(3):
            fn(i)
        if finish is not None:
            finish()
        barrier()
        k = min(args.steps, 100)
        t0 = time.perf_counter()
        e0.record()
        for i in range(k):
            fn(i)
        if finish is not None:
            finish()                                  # (asynchronous loss read-back: the last iteration's loss)
        e1.record()
        barrier()
        ms = max(e0.elapsed_time(e1), (time.perf_counter() - t0) * 1e3) / k
        if world > 1:
            t = torch.tensor([ms], device=dev, dtype=torch.float64)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            ms = float(t.item())
        return ms, k

    ms_dropin, k_e2e = time_e2e(e2e_step)
    ms_dropin_lag, _ = time_e2e(e2e_step_lagged, finish=e2e_lagged_finish)
    # the same call sites with the optimiser class swapped in create_optimizer (mneslam_b200.optim.Adam: one fused launch per tensor)
    ms_dropin_optim = ms_dropin_optim_lag = None
    if world == 1:
        try:
            from mneslam_b200.optim import Adam as MapAdam
            opt = MapAdam([{'params': model.decoder.parameters(), 'weight_decay': 1e-6, 'lr': mp['lr_decoder']},
                           {'params': model._flat_planes(), 'eps': 1e-15, 'lr': mp['lr_embed']}], betas=(0.9, 0.99))
            ms_dropin_optim, _ = time_e2e(e2e_step)
            ms_dropin_optim_lag, _ = time_e2e(e2e_step_lagged, finish=e2e_lagged_finish)
        except Exception as e:
            print('bench.py: drop-in with the fused optimiser class failed (%r)' % (e,), file=sys.stderr)
    # the fused driver shares the planes with the (now Parameter-wrapped) model: rebuild it on the same storage
    mapper2 = FusedMapper(model, world_size=world)
    mapper2.step_dev.copy_(mapper.step_dev)
    mapper = mapper2
    ms_e2e_eager, k_e2e = time_e2e(fused_e2e_step)
    ms_e2e, e2e_path = ms_e2e_eager, ('FusedMapper.step (C-ABI: sample, forward, losses, backward, Adam) per iteration, ray batch copied '
                                      'from pinned host memory, loss copied back to the host, no CUDA graph')
    try:
        from mneslam_b200.mapping import HostFedMapper
        hosted = HostFedMapper(mapper, db, n_rays).capture(host_rays[0], host_ids[0])
        ms_g, k_e2e = time_e2e(graphed_e2e_step)
        if ms_g < ms_e2e:
            ms_e2e, e2e_path = ms_g, ('HostFedMapper.step: host ray batch -> pinned staging -> ONE CUDA-graph launch (H2D copies, ray assembly, '
                                      'sample, forward, losses, backward, Adam, D2H of the loss) -> stream sync -> loss as a Python float; '
                                      'wall clock per iteration')
    except Exception as e:
        print('bench.py: host-fed graph unavailable (%r); e2e is the eager driver' % (e,), file=sys.stderr)
    ms_e2e_hostfed = ms_e2e
    ms_e2e_prefetch = None
    try:
        # data-loader style: each call trains on the previous call's batch and uploads / samples the next one under its optimiser
        from mneslam_b200.mapping import PrefetchingHostMapper
        pre = PrefetchingHostMapper(mapper, db, n_rays, spare_sms=args.spare_sms).prime(host_rays[0], host_ids[0])

        def prefetch_e2e_step(i):
            return pre.step(host_rays[(i + 1) % n_pool], host_ids[(i + 1) % n_pool])
        ms_p, k_e2e = time_e2e(prefetch_e2e_step)
        ms_e2e_prefetch = ms_p
        if ms_p < ms_e2e:
            ms_e2e, e2e_path = ms_p, ('PrefetchingHostMapper.step(next_batch): host ray batch -> pinned staging -> ONE CUDA-graph launch '
                                      '(iteration on the previous call\'s batch; H2D, ray assembly and sampling of this call\'s batch on a '
                                      'second stream under its optimiser; D2H of the loss) -> stream sync -> loss as a Python float; wall '
                                      'clock per iteration, one H2D batch and one D2H loss per call')
    except Exception as e:
        print('bench.py: prefetching host-fed graph unavailable (%r)' % (e,), file=sys.stderr)
    ms_e2e_async = None
    try:
        # the same, with the loss read back one call late too: the host never waits for the launch it has just made
        pre.close()
        pre2 = PrefetchingHostMapper(mapper, db, n_rays, spare_sms=args.spare_sms, async_loss=True).prime(host_rays[0], host_ids[0])

        def async_e2e_step(i):
            return pre2.step(host_rays[(i + 1) % n_pool], host_ids[(i + 1) % n_pool])
        ms_a, k_e2e = time_e2e(async_e2e_step, finish=pre2.flush)
        ms_e2e_async = ms_a
        pre2.close()
        if ms_a < ms_e2e:
            ms_e2e, e2e_path = ms_a, ('PrefetchingHostMapper(async_loss=True).step(next_batch): host ray batch -> pinned staging -> ONE '
                                      'CUDA-graph launch (iteration on the previous call\'s batch; H2D, ray assembly and sampling of this '
                                      'call\'s batch on a second stream under its optimiser; D2H of the loss) -> the call returns the loss '
                                      'of the iteration launched by the PREVIOUS call as a Python float (event wait), so the device runs '
                                      'the iterations back to back; the last loss is read before the timed region ends; wall clock per '
                                      'iteration, one H2D batch and one D2H loss per iteration')
    except Exception as e:
        print('bench.py: asynchronous-loss host-fed graph unavailable (%r)' % (e,), file=sys.stderr)
    e2e_value = pts_per_rank * world / (ms_e2e * 1e-3)

    # ---- tracking iteration (BASELINE config 3): ScanNet-shaped, 1024 rays x 117 samples, pose-only gradients
    tracking = None
    if world == 1:
        from mneslam_b200.tracking import PoseTracker
        tcfg = configs.scannet_scene0000_onegrid()
        torch.manual_seed(1)
        tmodel = JointEncoding(tcfg, torch.tensor(tcfg['mapping']['bound'], dtype=torch.float64)).to(dev)
        tb = {k: v.to(dev) for k, v in synthetic.make_ray_batch(tcfg, 1024, seed=3).items()}
        tracker = PoseTracker(tmodel, tb['c2w'], rot_rep='quat')
        for _ in range(5):
            tracker.step(tb['dirs_cam'], tb['target_rgb'], tb['target_d'])
        torch.cuda.synchronize()
        n_it = 20
        t0 = time.perf_counter()
        e0.record()
        for _ in range(n_it):
            tracker.step(tb['dirs_cam'], tb['target_rgb'], tb['target_d'])
        e1.record()
        torch.cuda.synchronize()
        ms_eager = max(e0.elapsed_time(e1), (time.perf_counter() - t0) * 1e3) / n_it
        ms_graph = None
        try:
            from mneslam_b200.tracking import GraphedPoseTracker
            gt = GraphedPoseTracker(tracker, tb['dirs_cam'], tb['target_rgb'], tb['target_d'])
            gt.run(3)
            torch.cuda.synchronize()
            t0 = time.perf_counter()
            e0.record()
            gt.run(n_it)
            e1.record()
            torch.cuda.synchronize()
            ms_graph = max(e0.elapsed_time(e1), (time.perf_counter() - t0) * 1e3) / n_it
        except Exception as e:
            print('bench.py: tracking graph capture failed (%r)' % (e,), file=sys.stderr)
        # the same iteration without autograd: mne_pose_rays + fused render + mne_pose_step (pose gradient and Adam in one kernel)
        ms_fused = ms_fused_eager = None
        try:
            from mneslam_b200.tracking import FusedTracker, GraphedPoseTracker
            ft = FusedTracker(tmodel, tb['c2w'], rot_rep='quat')
            for _ in range(5):
                ft.step(tb['dirs_cam'], tb['target_rgb'], tb['target_d'])
            torch.cuda.synchronize()
            t0 = time.perf_counter()
            e0.record()
            for _ in range(n_it):
                ft.step(tb['dirs_cam'], tb['target_rgb'], tb['target_d'])
            e1.record()
            torch.cuda.synchronize()
            ms_fused_eager = max(e0.elapsed_time(e1), (time.perf_counter() - t0) * 1e3) / n_it
            gf = GraphedPoseTracker(ft, tb['dirs_cam'], tb['target_rgb'], tb['target_d'])
            gf.run(3)
            torch.cuda.synchronize()
            t0 = time.perf_counter()
            e0.record()
            gf.run(n_it)
            e1.record()
            torch.cuda.synchronize()
            ms_fused = max(e0.elapsed_time(e1), (time.perf_counter() - t0) * 1e3) / n_it
        except Exception as e:
            print('bench.py: fused tracker failed (%r)' % (e,), file=sys.stderr)
        best = min(x for x in (ms_fused, ms_graph, ms_eager) if x is not None)
        tracking = {'ms_per_iter': best, 'iters': n_it,
                    'fused_tracker_graph': ms_fused, 'fused_tracker_eager': ms_fused_eager,
                    'autograd_tracker_graph': ms_graph, 'autograd_tracker_eager': ms_eager,
                    'workload': 'ScanNet-shaped (620x460, scene0000 bound, one grid): 1024 rays x 117 samples, map frozen, '
                                'quaternion + translation pose, Adam.  fused_tracker = FusedTracker.step (mne_pose_rays, sample, forward, '
                                'losses, ray-gradient backward, mne_pose_step; no autograd); autograd_tracker = PoseTracker.step (torch '
                                'autograd + torch.optim.Adam around the same fused render); graph = one CUDA-graph replay per iteration',
                    'points_per_iter': 1024 * 117}
        del tmodel, tracker

    inference = None
    if world == 1 and not args.no_extras:
        try:
            inference = inference_record(model, cfg, dev)
        except Exception as e:
            inference = {'error': repr(e)[:300]}
    gather_peak = None
    if rank == 0 and not args.no_extras:
        try:
            gather_peak = gather_peak_record(model, mapper, cfg, batch_db, n_rays, dev)
        except Exception as e:
            gather_peak = {'error': repr(e)[:300]}
    # ---- the records every N carries: configs[3] (strong scaling, hash grid) and the cross-rank self-check
    n_active_main = ops.backward_active_count(mapper._scratch, n_rays, s)
    extra = {}
    if not args.no_extras:
        try:
            extra['self_check'] = self_check_record(world, rank, dev)
        except Exception as e:
            extra['self_check'] = {'ok': False, 'error': repr(e)[:300]}
        if world > 1:
            try:
                extra['exchange'] = exchange_record(world, rank, dev)
            except Exception as e:
                extra['exchange'] = {'error': repr(e)[:300]}
        try:
            extra['configs3'] = configs3_record(world, rank, dev, args.steps)
        except Exception as e:
            extra['configs3'] = {'error': repr(e)[:300]}

    if rank == 0:
        peak, peak_src = measured_peak()
        # algorithmic bytes per launch (DESIGN.md "Roofline accounting"):
        #   forward : every sample gathers 3 planes x 2 levels x 4 texels x 128 B
        #   backward: the samples with a non-zero loss gradient scatter into the same texels (the others add nothing)
        #   Adam    : 4 B gradient read + 0.25 B of flags per parameter, 24 B (p, m, v in and out) per parameter of a group that
        #             ever received a gradient, 4 B (the gradient cleared) per parameter whose gradient was non-zero this step
        n_active = n_active_main
        fwd_bytes = pts_per_rank * ALG_BYTES_PER_POINT_FWD
        bwd_bytes = n_active * ALG_BYTES_PER_POINT_BWD
        n_param = mapper.total
        touched = touched_prof
        # every gradient group is read (4 B / parameter) with its flag (0.25 B); groups that ever received a gradient move p, m, v
        # in and out (24 B); the gradient is written back (cleared) only where it was non-zero this step (4 B)
        adam_bytes = int(n_param * (4 + (0.25 if mapper.touched is not None else 0.0) + touched * 24 + nz_frac * 4))
        if getattr(mapper, 'adam_dirty', False):
            # with the backward pass's texel flags (1 byte per 32 parameters) only the flagged gradient texels are read and cleared
            adam_bytes = int(n_param * (1 / 32 + 0.25 + touched * 24 + nz_frac * 8))
        gbs = lambda b, t_us: b / (t_us * 1e-6) / 1e9
        fwd_gbs, bwd_gbs, adam_gbs = gbs(fwd_bytes, us['fwd']), gbs(bwd_bytes, us['bwd']), gbs(adam_bytes, us['adam'])
        line = {
            'metric': METRIC, 'value': value, 'unit': UNIT, 'n_gpus': world, 'steps': args.steps, 'warmup': max(args.warmup, 3),
            'ms_per_step': ms_step, 'higher_is_better': True, 'scaling': 'weak', 'vs_baseline': None, 'dtype': 'f32',
            'data': 'synthetic', 'config': workload_config(cfg, n_rays, world),
            'cuda_graph': graph is not None, 'pipelined': pipe is not None,
            'gpu_launches': launches_per_step * args.steps,
            'e2e': {'value': e2e_value, 'unit': UNIT, 'h2d_bytes_per_step': h2d_bytes, 'd2h_bytes_per_step': 4,
                    'ms_per_step': ms_e2e, 'steps': k_e2e, 'path': e2e_path, 'ms_per_step_eager_driver': ms_e2e_eager,
                    'ms_per_step_host_fed_graph': ms_e2e_hostfed, 'ms_per_step_prefetching_host_fed_graph': ms_e2e_prefetch,
                    'ms_per_step_prefetching_async_loss': ms_e2e_async},
            'e2e_dropin': {'value': pts_per_rank * world / (ms_dropin * 1e-3), 'unit': UNIT, 'ms_per_step': ms_dropin,
                           'path': 'JointEncoding.forward -> loss.backward() -> torch.optim.Adam.step() (the reference call '
                                   'sites unchanged), same host buffers',
                           'ms_per_step_with_mneslam_b200_optim_Adam': ms_dropin_optim,
                           'loss_read_one_step_late': {'ms_per_step': ms_dropin_lag, 'ms_per_step_with_mneslam_b200_optim_Adam': ms_dropin_optim_lag,
                                                       'note': 'the same call sites; each call collects the loss of the PREVIOUS iteration '
                                                               '(pinned buffer + event), so the host never waits for the iteration it has just '
                                                               'queued -- the reference loop itself never reads the loss'}},
            'tracking': tracking,
            # the dominant kernel of the step is the optimiser; the two gather / scatter kernels follow
            'roofline': {'bound': 'hbm', 'kernel': 'adam_segments_kernel' if world == 1 else 'optimizer step: flag hand-over + barrier + adam_peers_sparse_kernel + barrier', 'achieved': adam_gbs, 'peak': peak, 'unit': 'GB/s',
                         'frac': adam_gbs / peak, 'traffic': ncu_traffic('adam_segments_kernel'), 'peak_source': peak_src,
                         'algorithmic_bytes_per_launch': adam_bytes, 'launch_us': us['adam'],
                         'touched_fraction': touched, 'nonzero_gradient_fraction': nz_frac,
                         'gradient_read_skipped_by_texel_flags': bool(getattr(mapper, 'adam_dirty', False)),
                         'points_fwd_kernel': {'achieved': fwd_gbs, 'frac': fwd_gbs / peak, 'launch_us': us['fwd'],
                                               'algorithmic_bytes_per_launch': fwd_bytes,
                                               'traffic': ncu_traffic('points_fwd_kernel'),
                                               'note': 'the fused render: encode, decode, compositing, loss sums, d loss / d raw + sample compaction in one launch'},
                         'points_bwd_kernel': {'achieved': bwd_gbs, 'frac': bwd_gbs / peak, 'launch_us': us['bwd'],
                                               'algorithmic_bytes_per_launch': bwd_bytes, 'active_samples': n_active,
                                               'traffic': ncu_traffic('points_bwd_kernel'),
                                               'note': 'points_bwd3_kernel: decoder backward + plane scatter over the active samples; the decoder weight gradients (breakdown_us.wgrad) run beside the optimiser'}},
            'breakdown_us': us,
            'fwd_bwd_points_per_s': pts_per_rank * world / ((us['sample'] + us['fwd'] + us['bwd']) * 1e-6),
            'clocks': clocks,
            'adam': {'exact_skip_of_never_touched_groups': mapper.touched is not None,
                     'touched_fraction_at_end': float(mapper.touched.float().mean().item()) if mapper.touched is not None else 1.0},
            'last_loss': float(mapper.total_loss(static_losses if graph is not None else losses).item()),
        }
        line.update(extra)
        if world > 1:
            # (the library's cross-rank barriers wait a bounded time: non-zero = a rank gave up waiting somewhere in this run)
            line['barrier_timeouts_at_end'] = mapper.barrier_timeouts()
        # the whole step against the gather roofline -- 6 144 algorithmic bytes per point (3 072 gathered + 3 072
        # of gradient contributions) x points/s over the measured HBM copy bandwidth of all GPUs
        line['roofline']['whole_step_frac_6144B_per_point'] = value * (ALG_BYTES_PER_POINT_FWD + ALG_BYTES_PER_POINT_BWD) / 1e9 / (peak * world)
        line['gather_peak'] = gather_peak
        line['inference'] = inference
        line['adam']['steady_state'] = steady
        if gather_peak and 'gather_gbs_after_l2_flush' in gather_peak:
            g_cold, g_warm = gather_peak['gather_gbs_after_l2_flush'], gather_peak['gather_gbs_warm']
            for k, gbs_k in (('points_fwd_kernel', fwd_gbs), ('points_bwd_kernel', bwd_gbs)):
                line['roofline'][k]['frac_l2'] = gbs_k / g_cold
                line['roofline'][k]['frac_l2_warm'] = gbs_k / g_warm
            line['roofline']['l2_gather_gbs'] = g_cold
            line['roofline']['l2_note'] = ('frac = algorithmic GB/s over the HBM copy bandwidth (MEASURED_PEAKS.json); frac_l2 = the same GB/s over '
                                           'the rate a loads-only kernel reaches on the SAME tap addresses after an L2 flush (gather_peak)')
        if world == 1 and not args.no_cpu_baseline:
            line['cpu_baseline'] = cpu_baseline_record(cfg, n_rays)
            try:
                del mapper, model, graph
                torch.cuda.empty_cache()
                line['torch_gpu_baseline'] = torch_gpu_record(cfg, n_rays, dev)
            except Exception as e:        # orientation only: never fail the bench line over it
                line['torch_gpu_baseline'] = {'unavailable': repr(e)[:200]}
        print(json.dumps(line), flush=True)
    if world > 1:
        # Tearing NCCL down while a captured graph still references its communicator hung the process once the JSON
        # line was out: synchronise, meet at a barrier, and leave without the destructor chain.
        torch.cuda.synchronize()
        dist.barrier()
        torch.cuda.synchronize()
        sys.stdout.flush()
        sys.stderr.flush()
        os._exit(0)


def main():
    args = parse()
    if args.impl == 'reference':
        run_reference(args)
    else:
        run_ours(args)


if __name__ == '__main__':
    main()
