#!/usr/bin/env python
"""Benchmark of the Co-SLAM render-and-optimise hot path (BASELINE.json metric: rays/sec of the fused
encode + MLP + render forward + backward).

    python bench.py --gpus N --steps K --warmup W                   # this repo's CUDA path
    python bench.py --impl reference --gpus N --steps K --warmup W  # the reference's CPU path (oracle port) on host cores

A step = one mapping-BA pass of the hot path over one ray batch: JointEncoding.forward (sampling, hash + OneBlob
encodings, SDF/colour decoders, volume rendering, four losses) and the backward of the weighted loss into the hash
table, the decoder and the ray origins/directions (the pose-gradient path).  N > 1: one process per GPU (torchrun),
rays sharded, losses / counts / gradients all-reduced over NCCL (coslam_b200/dist.py) -- weak scaling.
Prints ONE JSON line on rank 0.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

import torch                                   # noqa: E402

_STDOUT = sys.stdout                           # where the JSON line goes (see __main__)

from coslam_b200 import synth                  # noqa: E402

METRIC = 'rays_per_sec_render_fwd_bwd'
UNIT = 'rays/s'
BYTES_PER_POINT_FWD = 1024          # 16 levels x 8 corners x 8 B gathered           
BYTES_PER_POINT_BWD = 2048          # the same gathers again for dL/dx + 1024 B of gradient scattered to the table
# roofline.traffic is read from the newest ncu --set full summary under profiles/ (ncu_traffic_per_point), scaled by live points
NCU_V9_LIVE_POINTS = 6823377             # live points of the launch profiles/r1/ncu_full_v9_tc.txt captured (2^18 rays x 43 x 0.6053)


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument('--gpus', type=int, default=1)
    ap.add_argument('--steps', type=int, default=20)
    ap.add_argument('--warmup', type=int, default=5)
    ap.add_argument('--impl', default='ours', choices=['ours', 'reference'])
    ap.add_argument('--rays', type=int, default=1 << 22, help='rays per GPU per step (default: the top of BASELINE config 4\'s 2^16-2^22 sweep)')
    ap.add_argument('--quick', action='store_true', help='headline numbers only: no sweep, other configs, pipelines, strong scaling or GPU baseline')
    ap.add_argument('--workload', default='room0', choices=sorted(synth.CONFIGS))
    ap.add_argument('--hash-size', type=int, default=0, help='override grid.hash_size (scannet base file: 19)')
    ap.add_argument('--n-importance', type=int, default=-1, help='override training.n_importance (tum.yaml: 128)')
    ap.add_argument('--cpu-rays', type=int, default=4096, help='rays per CPU-baseline step')
    ap.add_argument('--no-cpu-baseline', action='store_true')
    ap.add_argument('--tracking', action='store_true', help='(default on one GPU) also time the tracking loop: ms/frame')
    ap.add_argument('--no-tracking', action='store_true')
    return ap.parse_args()


def loss_weights(cfg):
    t = cfg['training']
    return (t['rgb_weight'], t['depth_weight'], t['sdf_weight'], t['fs_weight'])


class ClockSampler(threading.Thread):
    """Samples SM clocks and throttle reasons of one GPU with nvidia-smi while the timed region runs."""
    Q = 'clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,' \
        'clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap'

    def __init__(self, index):
        super().__init__(daemon=True)
        self.index, self.rows, self.stop_flag, self.proc = index, [], False, None

    def run(self):
        try:
            self.proc = subprocess.Popen(['nvidia-smi', '-i', str(self.index), '--query-gpu=' + self.Q,
                                          '--format=csv,noheader,nounits', '-lms', '100'], stdout=subprocess.PIPE, text=True)
            for line in self.proc.stdout:
                if self.stop_flag:
                    break
                self.rows.append([c.strip() for c in line.split(',')])
        except Exception:
            pass

    def finish(self):
        self.stop_flag = True
        if self.proc is not None:
            self.proc.terminate()
        sm = sorted(int(r[0]) for r in self.rows if r and r[0].isdigit())
        mx = [int(r[1]) for r in self.rows if len(r) > 1 and r[1].isdigit()]
        names = ['hw_slowdown', 'hw_thermal_slowdown', 'sw_thermal_slowdown', 'sw_power_cap']
        reasons = [n for i, n in enumerate(names) if any(len(r) > 2 + i and r[2 + i].lower().startswith('active') for r in self.rows)]
        return {'sm_mhz': sm[len(sm) // 2] if sm else None, 'sm_max_mhz': max(mx) if mx else None, 'reasons': reasons,
                'samples': len(sm)}


def build_oracle(cfg, table, weights):
    from oracle.coslam_oracle import OracleScene
    orc = OracleScene(cfg, synth.bound_tensor(cfg))
    orc.load_weights(table, *weights)
    return orc


class ReferenceCpuModel(object):
    """The reference's OWN model/scene_rep.py:JointEncoding (unmodified copy under baseline/_ref, see
    __graft_entry__.install_reference) with the pure-torch tinycudann stand-in, on the host cores: kind "reference"."""
    kind = 'reference'

    def __init__(self, cfg, table, weights):
        from oracle import ref_loader
        cls = ref_loader.load_reference_model_class()
        self.cfg = cfg
        self.model = cls(cfg, synth.bound_tensor(cfg))
        with torch.no_grad():
            self.model.embed_fn.params.copy_(table)
            for dst, src in zip((self.model.decoder.sdf_net.model[0].weight, self.model.decoder.sdf_net.model[2].weight,
                                 self.model.decoder.color_net.model[0].weight, self.model.decoder.color_net.model[2].weight), weights):
                dst.copy_(src)
        self.model.train()

    def step(self, batch, noise):
        rays_o, rays_d, rgb, depth = batch
        for p in self.model.parameters():
            p.grad = None
        o = rays_o.clone().requires_grad_(True)
        d = rays_d.clone().requires_grad_(True)
        ret = self.model.forward(o, d, rgb, depth)                  # draws its own jitter (scene_rep.py:255)
        t = self.cfg['training']                                    # coslam.py:140-160
        loss = t['rgb_weight'] * ret['rgb_loss'] + t['depth_weight'] * ret['depth_loss'] + t['sdf_weight'] * ret['sdf_loss'] \
            + t['fs_weight'] * ret['fs_loss']
        loss.backward()
        return float(ret['rgb_loss'].detach())


class PortCpuModel(object):
    """oracle/coslam_oracle.py (the restatement, asserted equal to the reference in the build container): kind "port"."""
    kind = 'port'

    def __init__(self, cfg, table, weights):
        self.orc = build_oracle(cfg, table, weights)

    def step(self, batch, noise):
        return cpu_reference_step(self.orc, batch, noise)


def cpu_reference_step(orc, batch, noise):
    """One fwd+bwd of the reference path's CPU restatement (oracle/coslam_oracle.py == the reference's own torch code
    plus the pure-torch tinycudann stand-in), all host threads."""
    rays_o, rays_d, rgb, depth = batch
    for p in orc.parameters():
        p.grad = None
    o = rays_o.clone().requires_grad_(True)
    d = rays_d.clone().requires_grad_(True)
    r = orc.forward(o, d, rgb, depth, noise=noise)
    orc.total_loss(r).backward()
    return float(r['rgb_loss'].detach())


def time_cpu(cfg, n_rays, steps, warmup, budget_s=25.0, threads=None):
    """Returns (step times, threads used, kind): the reference's own CPU code when baseline/_ref holds it, else the port."""
    torch.set_num_threads(threads or os.cpu_count() or 1)
    weights = synth.decoder_init(seed=0)
    from oracle import ref_loader, tcnn_shim
    from oracle.coslam_oracle import hashgrid_encoding_config
    enc = hashgrid_encoding_config(cfg, synth.bound_tensor(cfg))
    meta = tcnn_shim.grid_meta(enc['n_levels'], enc['log2_hashmap_size'], enc['base_resolution'], enc['per_level_scale'])
    table = synth.table_init(meta['offset'][-1] * 2, 'trained')
    runner = (ReferenceCpuModel if ref_loader.available() else PortCpuModel)(cfg, table, weights)
    batch = synth.keyframe_ray_batch(cfg, n_rays, seed=0)[:4]
    S = cfg['training']['n_samples_d'] + cfg['training']['n_range_d']
    noise = torch.rand(n_rays, S)
    for _ in range(warmup):
        runner.step(batch, noise)
    times = []
    t_begin = time.perf_counter()
    for _ in range(steps):
        t0 = time.perf_counter()
        runner.step(batch, noise)
        times.append(time.perf_counter() - t0)
        if time.perf_counter() - t_begin > budget_s:
            break
    used = torch.get_num_threads()
    torch.set_num_threads(os.cpu_count() or 1)
    return times, used, runner.kind


CPU_KIND_TEXT = {'reference': "the reference's own model/scene_rep.py (unmodified copy in baseline/_ref) + pure-torch tinycudann shim",
                 'port': 'oracle port (reference scene_rep.py restated + pure-torch tinycudann shim)'}


def run_reference(args):
    """--impl reference: the reference's CPU implementation of the path on the box's host cores."""
    rank = int(os.environ.get('RANK', '0'))
    if rank != 0:
        return
    cfg = synth.CONFIGS[args.workload]()
    if args.hash_size:
        cfg['grid']['hash_size'] = args.hash_size
    if args.n_importance >= 0:
        cfg['training']['n_importance'] = args.n_importance
    steps = max(1, args.steps)
    times, cores, kind = time_cpu(cfg, args.cpu_rays, steps, max(1, min(args.warmup, 3)), budget_s=120.0)
    total = sum(times)
    value = args.cpu_rays * len(times) / total
    S = cfg['training']['n_samples_d'] + cfg['training']['n_range_d']
    sample = '%d steps of %d rays x %d samples, fwd+bwd, %s' % (len(times), args.cpu_rays, S, CPU_KIND_TEXT[kind])
    print(json.dumps({
        'impl': 'reference', 'metric': METRIC, 'value': value, 'unit': UNIT, 'n_gpus': args.gpus, 'steps': len(times),
        'warmup': args.warmup, 'ms_per_step': 1e3 * total / len(times), 'higher_is_better': True, 'scaling': 'weak',
        'vs_baseline': None, 'dtype': 'f32', 'data': 'synthetic',
        'config': {'workload': '%s mapping-BA render fwd+bwd, %d-ray CPU sample of the %d-ray/GPU batch' % (args.workload, args.cpu_rays, args.rays),
                   'rays_per_step': args.cpu_rays, 'samples_per_ray': S},
        'cpu_baseline': {'value': value, 'unit': UNIT, 'cores': cores, 'kind': kind, 'sample': sample},
        'e2e': {'value': value, 'unit': UNIT, 'h2d_bytes_per_step': 0, 'd2h_bytes_per_step': 0},
        'gpu_launches': 0}), file=_STDOUT, flush=True)


def run_ours(args):
    import torch.distributed as dist
    from coslam_b200 import _lib, ops
    from coslam_b200.dist import ShardedMapper
    from coslam_b200.scene_rep import JointEncoding

    world = int(os.environ.get('WORLD_SIZE', '1'))
    rank = int(os.environ.get('RANK', '0'))
    local = int(os.environ.get('LOCAL_RANK', '0'))
    if not torch.cuda.is_available():
        raise SystemExit('bench.py: no CUDA device (the product path has no CPU fallback; use --impl reference for the CPU arm)')
    torch.cuda.set_device(local)
    dev = torch.device('cuda', local)
    if world > 1:
        dist.init_process_group('nccl', device_id=dev)

    cfg = synth.CONFIGS[args.workload]()
    if args.hash_size:
        cfg['grid']['hash_size'] = args.hash_size
    if args.n_importance >= 0:
        cfg['training']['n_importance'] = args.n_importance
    S = cfg['training']['n_samples_d'] + cfg['training']['n_range_d']
    n = args.rays
    model = JointEncoding(cfg, synth.bound_tensor(cfg).to(dev)).to(dev)
    model.noise_mode = 'device'                      # in-kernel Philox jitter (throughput mode; parity tests use 'host')
    model.noise_seed = 1 + rank
    with torch.no_grad():
        model.embed_fn.params.copy_(synth.table_init(model.embed_fn.params.numel(), 'trained').to(dev))
        for dst, src in zip(model.decoder.fused_weights(), synth.decoder_init(seed=0)):
            dst.copy_(src.to(dev))
    model.train()
    host = [t.contiguous().pin_memory() for t in synth.keyframe_ray_batch(cfg, n, seed=100 + rank)[:4]]
    resident = [t.to(dev) for t in host]
    mapper = ShardedMapper(model, loss_weights(cfg))

    def step(batch):
        for p in model.parameters():
            p.grad = None
        o = batch[0].requires_grad_(True)
        d = batch[1].requires_grad_(True)
        o.grad = d.grad = None
        loss, _ = mapper.step(o, d, batch[2], batch[3])
        return loss

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def timed(fn, steps):
        barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(steps):
            fn()
        e1.record()
        barrier()
        ms = torch.tensor([e0.elapsed_time(e1)], device=dev)
        if world > 1:
            dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        return float(ms)

    # ---- device-resident throughput ------------------------------------------------------------------------------
    for _ in range(max(3, args.warmup)):
        step(resident)
    sampler = ClockSampler(local)
    sampler.start()
    time.sleep(0.3)
    ops.KERNEL_EVENTS = []
    launches0 = _lib.KERNEL_LAUNCHES
    ms = timed(lambda: step(resident), args.steps)
    launches = _lib.KERNEL_LAUNCHES - launches0
    events, ops.KERNEL_EVENTS = ops.KERNEL_EVENTS, None
    clocks = sampler.finish()
    kt = {}
    for name, a, b in events:
        kt.setdefault(name, []).append(a.elapsed_time(b))
    k_fwd = sum(kt['points_fwd']) / args.steps          # per step: the lazy forward launches the per-point kernel twice
    k_bwd = sum(kt['points_bwd']) / args.steps
    value = world * n * args.steps / (ms * 1e-3)
    # fraction of points with a non-zero output gradient (the backward skips the rest: exact zeros), one untimed step
    ops.KEEP_LAST_D_RAW = ops.KEEP_LAST_MASK = True
    step(resident)
    live_frac = float((ops.LAST_D_RAW.reshape(-1, 4) != 0).any(1).float().mean())
    eval_frac = float(ops.LAST_EVALUATED.float().mean()) if ops.LAST_EVALUATED is not None else 1.0
    ops.KEEP_LAST_D_RAW, ops.LAST_D_RAW, ops.KEEP_LAST_MASK, ops.LAST_MASK, ops.LAST_EVALUATED = False, None, False, None, None
    del ops.MASK_LOG[:]

    # ---- end to end through the public API: pinned host rays in, loss out, every step -----------------------------
    # Every step copies its rays from pinned host memory and reads its loss back.  The copy of step k + 1 is issued on a copy stream
    # in front of step k's launches (two staging sets), as a training loop that prefetches its next batch does: one copy per step inside
    # the timed region, overlapped with the previous step's kernels instead of in front of its own.
    staging = [[torch.empty_like(t, device=dev) for t in host] for _ in range(2)]
    copy_stream = torch.cuda.Stream(device=dev)
    copied = [torch.cuda.Event(), torch.cuda.Event()]
    slot = [0]

    def prefetch(i):
        with torch.cuda.stream(copy_stream):
            for dst, src in zip(staging[i], host):
                dst.copy_(src, non_blocking=True)
            copied[i].record(copy_stream)

    prefetch(0)

    def e2e_step():
        cur = slot[0]
        slot[0] = cur ^ 1
        # staging[cur ^ 1] is free: the step that read it ended with a read-back of its loss
        prefetch(cur ^ 1)
        torch.cuda.current_stream().wait_event(copied[cur])
        st = staging[cur]
        loss = step([st[0].detach(), st[1].detach(), st[2], st[3]])
        return float(loss.detach().cpu())            # device -> host read of the step's result

    for _ in range(3):
        e2e_step()
    ms_e2e = timed(e2e_step, args.steps)
    e2e_value = world * n * args.steps / (ms_e2e * 1e-3)
    h2d = sum(t.numel() * t.element_size() for t in host)

    extra = {}
    full = not args.quick
    if full:
        # ---- the default mode of the drop-in (COSLAM_B200_NOISE=host): the stream of CPU torch.rand(N, S) inside every step ----
        # (at this size generated on the device from host-walked Mersenne-Twister checkpoints: ops.device_uniform)
        model.noise_mode = 'host'
        e2e_step()
        k = min(3, args.steps)
        ms_host = timed(e2e_step, k)
        model.noise_mode = 'device'
        extra['e2e_host_noise'] = {'value': world * n * k / (ms_host * 1e-3), 'unit': UNIT, 'ms_per_step': ms_host / k, 'steps': k,
                                   'h2d_bytes_per_step': h2d + (n * S * 4 if n * S < ops.DEVICE_STREAM_MIN else (n * S // (624 * ops.DEVICE_STREAM_SEGMENT) + 1) * 2496),
                                   'd2h_bytes_per_step': 4,
                                   'note': "the reference's own jitter stream, torch.rand(z_vals.shape) of the CPU generator (scene_rep.py:255), bit for bit: "
                                           "batches of >= 2^20 floats are generated on the device from Mersenne-Twister checkpoints the host walks "
                                           "(the generator state advances as if torch.rand had run); smaller ones on the host + pinned upload"}
        # ---- BASELINE config 4's ray sweep, device-resident, same step -----------------------------------------------------
        sweep = {}
        for lg in (16, 18, 20, 22):
            if (1 << lg) == n:
                sweep['2^%d' % lg] = {'rays_per_gpu': n, 'ms_per_step': ms / args.steps, 'value': value, 'unit': UNIT}
                continue
            b = [t.to(dev) for t in synth.keyframe_ray_batch(cfg, 1 << lg, seed=100 + rank)[:4]]
            for _ in range(3):
                step(b)
            k = 5 if lg >= 20 else 20
            t_ms = timed(lambda: step(b), k)
            sweep['2^%d' % lg] = {'rays_per_gpu': 1 << lg, 'ms_per_step': t_ms / k, 'value': world * (1 << lg) * k / (t_ms * 1e-3), 'unit': UNIT}
            del b
        extra['sweep'] = sweep
        # ---- the same step with FIXED rays (no pose is optimised: first-frame mapping, mapping on given poses): NOT the headline ----
        # the backward then reads the forward's kept hash-grid features instead of gathering the table again (cb200_points_bwd_kept)
        from coslam_b200 import ops
        lg = min(20, n.bit_length() - 1)
        b = [t.to(dev) for t in synth.keyframe_ray_batch(cfg, 1 << lg, seed=100 + rank)[:4]]

        def fixed_step():
            for p in model.parameters():
                p.grad = None
            return mapper.step(b[0], b[1], b[2], b[3])[0]
        mo = {'rays_per_gpu': 1 << lg, 'unit': UNIT, 'note': 'rays without gradient; the headline step also produces the ray (pose) gradients'}
        for keep in (True, False):
            ops.KEEP_FEATURES = keep
            for _ in range(3):
                fixed_step()
            t_ms = timed(fixed_step, 5)
            mo['kept_features' if keep else 'gathering_backward'] = {'ms_per_step': t_ms / 5, 'value': world * (1 << lg) * 5 / (t_ms * 1e-3)}
        ops.KEEP_FEATURES = True
        extra['map_only'] = mo
        del b
        for p in model.parameters():
            p.grad = None
        torch.cuda.empty_cache()
        group = dist.group.WORLD if world > 1 else None
        try:
            extra['strong_scaling'] = strong_scaling(dev, group, world, rank)
        except Exception as e:
            extra['strong_scaling'] = {'error': '%s: %s' % (type(e).__name__, e)}
    gather_peak = None
    if rank == 0 and full:
        try:
            gather_peak = measure_gather_peak(model, cfg, dev)
        except Exception as e:
            gather_peak = {'error': '%s: %s' % (type(e).__name__, e)}
    if rank == 0 and (args.tracking or world == 1) and not args.no_tracking:
        extra['tracking'] = time_tracking(model, cfg, dev)
        if cfg['grid']['oneGrid']:
            extra['mapping'] = time_mapping(model, cfg, dev)
        if full:
            extra['other_configs'] = other_config_loops(dev)
            extra['pipelines'] = pipelines()
            try:
                extra['gpu_torch_baseline'] = gpu_torch_baseline(cfg, dev)
            except Exception as e:
                extra['gpu_torch_baseline'] = {'error': '%s: %s' % (type(e).__name__, e)}

    cpu = None
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        times, cores, kind = time_cpu(cfg, args.cpu_rays, steps=1000, warmup=1, budget_s=12.0)
        cpu = {'value': args.cpu_rays * len(times) / sum(times), 'unit': UNIT, 'cores': cores, 'kind': kind,
               'sample': '%d steps of %d rays x %d samples fwd+bwd (%s, torch CPU, %d threads)' % (len(times), args.cpu_rays, S, CPU_KIND_TEXT[kind], cores)}
        t1, _, _ = time_cpu(cfg, args.cpu_rays, steps=1000, warmup=1, budget_s=8.0, threads=1)       # the per-core figure
        cpu['one_thread'] = {'value': args.cpu_rays * len(t1) / sum(t1), 'unit': UNIT, 'cores': 1, 'steps': len(t1)}

    if rank == 0:
        peaks = {}
        try:
            peaks = json.load(open(os.path.join(ROOT, 'MEASURED_PEAKS.json')))
        except Exception:
            pass
        peak = float(peaks.get('hbm_gbs', 6650.0))
        P = n * S
        tc = os.environ.get('COSLAM_B200_DECODER', 'tc') != 'simt'
        # algorithmic bytes of the backward: gathers + scatter only for the points that carry a gradient
        bwd_bytes = P * live_frac * BYTES_PER_POINT_BWD
        achieved = bwd_bytes / (k_bwd * 1e-3) / 1e9
        live_points = P * live_frac
        traffic_src = ncu_traffic_per_point('points_bwd_tc_kernel') if tc else None
        traffic = None
        if traffic_src is not None:
            # the capture's DRAM bytes scale with the live points of a launch (the table stays in L2; HBM sees the per-point streams)
            per_point = traffic_src['dram_bytes_per_launch'] / (traffic_src['live_points'] or NCU_V9_LIVE_POINTS)
            traffic = per_point * live_points
        from coslam_b200 import ops as ops_mod
        roofline = {'bound': 'hbm', 'kernel': 'points_bwd_tc_kernel<wgrad>' if tc else 'points_bwd_kernel<wgrad>', 'achieved': achieved, 'peak': peak, 'unit': 'GB/s',
                    'frac': achieved / peak, 'traffic': traffic,
                    'traffic_source': ('%s: dram bytes per live point of that capture x live points of this launch' % traffic_src['file']) if traffic_src else None,
                    'peak_source': 'measured (MEASURED_PEAKS.json hbm_gbs)' if peaks else 'fallback',
                    'algorithmic_bytes_per_launch': bwd_bytes, 'live_point_fraction': live_frac, 'avg_launch_ms': k_bwd,
                    'fwd_kernel': {'kernel': 'points_fwd_tc_kernel' if tc else 'points_fwd_kernel', 'avg_launch_ms': k_fwd,
                                   'evaluated_point_fraction': eval_frac,
                                   'achieved': P * eval_frac * BYTES_PER_POINT_FWD / (k_fwd * 1e-3) / 1e9},
                    # the training step keeps rows in the forward so that the backward does not gather twice (ops.KEEP_FEATURES /
                    # KEEP_DERIVATIVES): the backward's fraction rose at the forward's expense, so the pair is stated as well
                    'kept_rows': {'enabled': bool(tc and ops_mod.KEEP_FEATURES and ops_mod.KEEP_DERIVATIVES), 'bytes_per_point': 512,
                                  'note': 'same algorithmic bytes (SURVEY 8d: 1 024 B/pt forward, 2 048 B/pt backward); the backward streams 512 B/pt '
                                          'written by the forward instead of gathering 1 024 B/pt again'},
                    'fwd_bwd_pair': {'achieved': (bwd_bytes + P * eval_frac * BYTES_PER_POINT_FWD) / ((k_fwd + k_bwd) * 1e-3) / 1e9,
                                     'frac': (bwd_bytes + P * eval_frac * BYTES_PER_POINT_FWD) / ((k_fwd + k_bwd) * 1e-3) / 1e9 / peak},
                    'share_of_step': (k_fwd + k_bwd) / (ms / args.steps),
                    'kernel_ms_per_step': {k: sum(v) / args.steps for k, v in kt.items()}}
        if gather_peak is not None and 'gather_reduce_GBps' in gather_peak:
            # the physical ceiling: the table is L2-resident, so the kernels are bounded by the L2 / L1TEX gather + reduction rate at
            # this access pattern, measured live with no decoder at all (state both)
            roofline['l2_gather_peak'] = gather_peak
            roofline['frac_l2_gather'] = achieved / gather_peak['gather_reduce_GBps']
            roofline['fwd_kernel']['frac_l2_gather'] = roofline['fwd_kernel']['achieved'] / gather_peak['gather_GBps']
        elif gather_peak is not None:
            roofline['l2_gather_peak'] = gather_peak
        out = {
            'metric': METRIC, 'value': value, 'unit': UNIT, 'n_gpus': world, 'steps': args.steps, 'warmup': max(3, args.warmup),
            'ms_per_step': ms / args.steps, 'higher_is_better': True, 'scaling': 'weak', 'vs_baseline': None, 'dtype': 'f32',
            'data': 'synthetic',
            'config': {'workload': '%s mapping-BA render fwd+bwd (hash 2^%d, 16x2 levels, OneBlob 16, MLP 80-32-16/63-32-3), %d rays/GPU x %d samples'
                       % (args.workload, cfg['grid']['hash_size'], n, S),
                       'rays_per_gpu': n, 'samples_per_ray': S, 'noise': 'device philox',
                       'l2': 'per-step intermediates (z, raw, d_raw, masks: %.0f MB) exceed the 126 MB L2; the %.1f MB hash table is L2-resident by design'
                       % (P * 44 / 1e6, model.embed_fn.params.numel() * 4 / 1e6)},
            'clocks': clocks,
            'e2e': {'value': e2e_value, 'unit': UNIT, 'h2d_bytes_per_step': h2d, 'd2h_bytes_per_step': 4, 'ms_per_step': ms_e2e / args.steps,
                    'copies': 'one H2D copy of the step\'s rays from pinned memory and one D2H read of its loss per step, inside the timed region; '
                              'the copy of step k+1 runs on a copy stream while step k computes (two staging sets)'},
            'gpu_launches': launches,
            'roofline': roofline,
            'cpu_baseline': cpu,
        }
        out.update(extra)
        print(json.dumps(out), file=_STDOUT, flush=True)
    if world > 1:
        dist.destroy_process_group()


# ------------------------------------------------------------------------------------------------------------------------
# extra measurements of the line (VERDICT r1 "Next round" items 2 and 3)
# ------------------------------------------------------------------------------------------------------------------------

def measure_gather_peak(model, cfg, dev, n_rays=49152):
    """Measured L2 gather / reduction ceiling at the fused kernels' access pattern for THIS table (tools/l2_gather_peak.py,
    csrc/microbench.cu): GB/s of table bytes with no decoder and no index arithmetic -- the physical denominator next to the
    HBM figure (the table is L2-resident, SURVEY.md 8d: "state both")."""
    import ctypes
    from coslam_b200 import _lib, ops
    sys.path.insert(0, os.path.join(ROOT, 'tools'))
    import l2_gather_peak as lg
    lg.DEV = dev
    rays_o, rays_d, _, depth = synth.ray_batch(cfg, n_rays, seed=0)
    z = ops.sample_rays(model.spec, depth.to(dev), perturb=False)
    n, s = z.shape
    pts = torch.empty((n * s, 3), device=dev)
    x = torch.empty((n * s, 3), device=dev)
    o, d = rays_o.to(dev).contiguous(), rays_d.to(dev).contiguous()
    _lib.check(_lib.load().cb200_debug_positions(ctypes.byref(model.spec.scene()), _lib.ptr(o), _lib.ptr(d), _lib.ptr(z), ctypes.c_int64(n),
                                                 ctypes.c_int64(s), _lib.ptr(pts), _lib.ptr(x), _lib.stream()), 'debug_positions')
    idx = lg.corner_list(model, x)
    table = model.embed_fn.params.detach()
    dtable = torch.zeros_like(table)
    P = idx.shape[0]
    out = {'points': P, 'pattern': 'corner list of %d rays x %d samples of this workload, lane pair per point, no other arithmetic' % (n, s)}
    for mode, key, passes in ((1, 'gather', 1), (2, 'reduce', 1), (3, 'gather_reduce', 2)):
        ms = lg.time_mode(model, idx, table, dtable, mode, iters=10, warmup=3)
        out[key + '_GBps'] = P * 1024.0 * passes / (ms * 1e-3) / 1e9
    return out


def ncu_traffic_per_point(kernel='points_bwd_tc_kernel'):
    """dram__bytes_read.sum + dram__bytes_write.sum of the newest committed `ncu --set full` summary of `kernel` under
    profiles/ (tools/ncu_summary.py format), per LIVE point of that capture when the file states it, else per launch."""
    import glob
    import re
    best = None
    for path in sorted(glob.glob(os.path.join(ROOT, 'profiles', 'r*', 'ncu_full_*.txt'))):
        text = open(path).read()
        if kernel not in text:
            continue
        pts = re.search(r'live_points:\s*([0-9]+)', text)
        # the summary may hold several kernels: read the section of the first launch of `kernel`
        sec = [b for b in text.split('\nkernel: ')[1:] if kernel in b.split('\n', 1)[0]]
        if not sec:
            continue
        text = sec[0]
        rd = re.search(r'dram__bytes_read\.sum\s+([0-9.]+) (\w)byte', text)
        wr = re.search(r'dram__bytes_write\.sum\s+([0-9.]+) (\w)byte', text)
        if not (rd and wr):
            continue
        unit = {'K': 1e3, 'M': 1e6, 'G': 1e9}
        total = float(rd.group(1)) * unit[rd.group(2)] + float(wr.group(1)) * unit[wr.group(2)]
        rnd = int(re.search(r'profiles/r(\d+)/', path.replace(os.sep, '/')).group(1))
        ver = re.search(r'_v(\d+)', os.path.basename(path))
        key = (rnd, int(ver.group(1)) if ver else 0)
        if best is None or key > best[0]:
            best = (key, path, total, int(pts.group(1)) if pts else None)
    if best is None:
        return None
    _, path, total, live = best
    return {'file': os.path.relpath(path, ROOT), 'dram_bytes_per_launch': total, 'live_points': live}


def gpu_torch_baseline(cfg, dev, sizes=(4096, 65536)):
    """The closest stand-in for the reference's GPU path this box allows (tinycudann is not installed): the reference's model
    code restated (oracle/coslam_oracle.py == model/scene_rep.py + decoder.py, asserted equal in the build container) with the
    pure-torch HashGrid / OneBlob shim, moved to the same B200 -- ATen + cuBLAS fp32, fwd + bwd of the weighted loss, CUDA-synced
    median.  Checker code timed as a BASELINE only (like cpu_baseline); never on the product path."""
    from oracle.coslam_oracle import OracleScene
    from oracle import tcnn_shim
    from oracle.coslam_oracle import hashgrid_encoding_config
    enc = hashgrid_encoding_config(cfg, synth.bound_tensor(cfg))
    meta = tcnn_shim.grid_meta(enc['n_levels'], enc['log2_hashmap_size'], enc['base_resolution'], enc['per_level_scale'])
    orc = OracleScene(cfg, synth.bound_tensor(cfg))
    orc.load_weights(synth.table_init(meta['offset'][-1] * 2, 'trained'), *synth.decoder_init(seed=0))
    orc.to(dev)
    S = cfg['training']['n_samples_d'] + cfg['training']['n_range_d']
    out = {'what': 'oracle restatement of the reference model + pure-torch tinycudann shim on cuda (ATen / cuBLAS fp32), fwd+bwd, median',
           'caveat': 'tinycudann itself is not installed on this box; its fused CUDA encodings are faster than this stand-in'}
    for n in sizes:
        batch = [t.to(dev) for t in synth.keyframe_ray_batch(cfg, n, seed=0)[:4]]
        noise = torch.rand(n, S, device=dev)
        times = []
        for it in range(2 + 5):
            for p in orc.parameters():
                p.grad = None
            o = batch[0].clone().requires_grad_(True)
            d = batch[1].clone().requires_grad_(True)
            torch.cuda.synchronize()
            t0 = time.perf_counter()
            r = orc.forward(o, d, batch[2], batch[3], noise=noise)
            orc.total_loss(r).backward()
            torch.cuda.synchronize()
            if it >= 2:
                times.append(time.perf_counter() - t0)
        med = sorted(times)[len(times) // 2]
        out['rays_%d' % n] = {'ms_per_step': med * 1e3, 'value': n / med, 'unit': UNIT}
        del batch, noise, r
        torch.cuda.empty_cache()
    return out



def other_config_loops(dev):
    """tracking ms/frame and mapping ms/iteration for BASELINE configs 3 and 5's shapes (scene0000: S = 117, quaternion poses,
    lattice 63^3; fr1_desk: S = 85 + 128 importance samples, 20 BA iterations, lattice 63^3), random-init maps."""
    from coslam_b200.scene_rep import JointEncoding
    out = {}
    for name, cfg in (('scannet0000', synth.scannet0000_config()), ('tum_desk', synth.tum_desk_config(n_importance=128))):
        try:
            if name == 'tum_desk':            # the frames fr1_desk.yaml:18-19 produces after resize + crop
                cfg['cam'].update({'H': 368, 'W': 496, 'fx': 413.84, 'fy': 413.2, 'cx': 246.88, 'cy': 196.24})
            model = JointEncoding(cfg, synth.bound_tensor(cfg).to(dev)).to(dev)
            model.noise_mode = 'device'
            with torch.no_grad():
                model.embed_fn.params.copy_(synth.table_init(model.embed_fn.params.numel(), 'trained').to(dev))
                for dst, src in zip(model.decoder.fused_weights(), synth.decoder_init(seed=0)):
                    dst.copy_(src.to(dev))
            model.train()
            out[name] = {'tracking': time_tracking(model, cfg, dev, frames=10, full=False), 'mapping': time_mapping(model, cfg, dev, calls=3)}
            del model
        except Exception as e:                                   # a failure here must not take the headline line with it
            out[name] = {'error': '%s: %s' % (type(e).__name__, e)}
        torch.cuda.empty_cache()
    return out


def pipelines(frames=40):
    """BASELINE configs 2, 3 and 5 in miniature (tools/slam_synth.py: sphere-traced synthetic RGB-D frames of the named
    resolutions, first-frame mapping, per-frame tracking, BA every 5th frame, keyframe store): steady-state frames/s, tracking
    ms/frame, mapping ms/iteration and the ATE against the generator's poses."""
    sys.path.insert(0, os.path.join(ROOT, 'tools'))
    import slam_synth
    out = {}
    for name in ('room0', 'scannet0000', 'tum_desk'):
        try:
            out[name] = slam_synth.run_workload(name, frames=frames)
        except Exception as e:
            out[name] = {'error': '%s: %s' % (type(e).__name__, e)}
        torch.cuda.empty_cache()
    return out


def strong_scaling(dev, group, world, rank, log2s=(18, 22), iters=5, n_kf=401):
    """BASELINE config 4 as worded: ONE batch of N rays per BA iteration sharded over the ranks (N / world each) through
    coslam_b200.mapping.BundleAdjuster(group=...): ray assembly over a 401-keyframe device store, forward, losses, backward
    (table, decoder, 400 poses), smoothness, Adam; counts / loss sums and the gradient bucket all-reduced over NCCL.  Device time,
    max over ranks.  At world = 1 this is the denominator of the strong-scaling efficiency."""
    import random
    import torch.distributed as dist
    from coslam_b200 import optim as cb_optim
    from coslam_b200 import pose
    from coslam_b200.keyframe import KeyFrameDatabase
    from coslam_b200.mapping import BundleAdjuster
    from coslam_b200.scene_rep import JointEncoding
    cfg = synth.room0_config()
    m, cam = cfg['mapping'], cfg['cam']
    H, W = cam['H'], cam['W']
    R = int(H * W * m['n_pixels'])
    torch.manual_seed(0)
    random.seed(0)
    model = JointEncoding(cfg, synth.bound_tensor(cfg).to(dev)).to(dev)
    with torch.no_grad():
        model.embed_fn.params.copy_(synth.table_init(model.embed_fn.params.numel(), 'trained', seed=77).to(dev))
        for dst, src in zip(model.decoder.fused_weights(), synth.decoder_init(seed=5)):
            dst.copy_(src.to(dev))
    model.noise_mode = 'device'
    model.train()
    db = KeyFrameDatabase(cfg, H, W, n_kf, R, dev)
    g = torch.Generator(device=dev).manual_seed(1)
    dirs = synth.camera_ray_dirs(H, W, cam['fx'], cam['fy'], cam['cx'], cam['cy']).reshape(-1, 3).to(dev)
    pick = torch.randint(0, H * W, (n_kf, R), device=dev, generator=g)
    db.rays[:n_kf, :, :3] = dirs[pick]
    db.rays[:n_kf, :, 3:6] = torch.rand((n_kf, R, 3), device=dev, generator=g)
    db.rays[:n_kf, :, 6] = torch.rand((n_kf, R), device=dev, generator=g) * 3.0 + 0.5
    db.frame_ids = torch.arange(n_kf) * m['keyframe_every']
    bound = synth.bound_tensor(cfg)
    centre, ext = bound.mean(1), bound[:, 1] - bound[:, 0]
    gh = torch.Generator().manual_seed(2)
    poses = []
    for k in range(n_kf + 1):
        r = torch.randn(1, 3, generator=gh)
        poses.append(pose.pose_matrix(r / r.norm() * (0.2 + 0.001 * k), (centre + 0.1 * ext * (torch.rand(3, generator=gh) - 0.5))[None])[0])
    kf_poses, cur_pose = torch.stack(poses[:n_kf]).to(dev), poses[n_kf].to(dev)
    current = torch.cat([dirs, torch.rand((H * W, 3), device=dev, generator=g), torch.rand((H * W, 1), device=dev, generator=g) * 3 + 0.5], -1)
    opt = cb_optim.Adam([{'params': model.decoder.parameters(), 'weight_decay': 1e-6, 'lr': m['lr_decoder']},
                         {'params': model.embed_fn.parameters(), 'eps': 1e-15, 'lr': m['lr_embed']}], betas=(0.9, 0.99))
    ba = BundleAdjuster(model, cfg, db, opt, dev, group=group, device_sampler=True)
    m['iters'] = iters
    out = {'scaling': 'strong', 'keyframes': n_kf, 'iters_per_call': iters}
    for lg in log2s:
        m['sample'] = 1 << lg
        m['min_pixels_cur'] = max(100, (1 << lg) // 64)
        ba.run(kf_poses, cur_pose, current)
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        reps = 3
        for _ in range(reps):
            ba.run(kf_poses, cur_pose, current)
        e1.record()
        torch.cuda.synchronize()
        ms = torch.tensor([e0.elapsed_time(e1) / (reps * iters)], device=dev)
        if world > 1:
            dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        n_rays = m['sample'] + max(m['sample'] // n_kf, m['min_pixels_cur'])
        out['rays_2^%d' % lg] = {'rays_per_iteration': n_rays, 'ms_per_iteration': float(ms), 'value': n_rays / (float(ms) * 1e-3), 'unit': UNIT}
    del ba, db, model
    torch.cuda.empty_cache()
    return out



def time_tracking(model, cfg, dev, frames=20, full=True):
    """Tracking ms/frame (BASELINE.json metric 2): `tracking.iter` dependent iterations of pose -> rays -> fwd -> loss -> bwd to
    the pose -> pose Adam, with the best-pose bookkeeping (coslam.py:502-573), through coslam_b200.tracking.FrameTracker:
    the loop as ONE CUDA graph per frame, and the same loop launched eagerly for comparison.  Inputs start on the host
    (pinned), the estimated pose is read back every frame."""
    from coslam_b200 import pose
    from coslam_b200.tracking import FrameTracker
    import random
    from coslam_b200.keyframe import sample_range
    n_rays = int(cfg['tracking']['sample'])
    g = torch.Generator().manual_seed(5)
    bound = synth.bound_tensor(cfg)
    cam, trk = cfg['cam'], cfg['tracking']
    H, W, iH, iW = cam['H'], cam['W'], trk['ignore_edge_H'], trk['ignore_edge_W']
    # one synthetic RGB-D frame on the HOST, as the reference's DataLoader serves it (datasets/dataset.py:250-265)
    frame_dirs = synth.camera_ray_dirs(H, W, cam['fx'], cam['fy'], cam['cx'], cam['cy'])[iH:H - iH, iW:W - iW, :].contiguous()
    frame_rgb = torch.rand(H - 2 * iH, W - 2 * iW, 3, generator=g)
    frame_depth = torch.rand(H - 2 * iH, W - 2 * iW, generator=g) * 3 + 0.5
    Hh = H - 2 * iH
    random.seed(11)

    def draw():
        """coslam.py:537-544: `tracking.sample` pixels by random.sample (bit-identical draws in C) and the CPU image gather."""
        idx = torch.from_numpy(sample_range(Hh * (W - 2 * iW), n_rays))
        ih, iw = idx % Hh, idx // Hh
        return frame_dirs[ih, iw, :], frame_rgb[ih, iw, :], frame_depth[ih, iw].unsqueeze(-1)

    c2w0 = pose.pose_matrix(torch.tensor([[0.3, -0.5, 0.2]]), bound.mean(1)[None])[0]
    out = {'rays': n_rays, 'iters': int(cfg['tracking']['iter']), 'frame': '%dx%d' % (W, H),
           'timed_region': 'per frame: random.sample pixel draw + CPU gather of (direction, rgb, depth) + upload + %d iterations + pose read-back'
                           % int(cfg['tracking']['iter'])}
    noise_state = (model.noise_mode, model._noise_counter)
    fusable = bool(cfg['grid']['oneGrid']) and cfg['training']['n_importance'] == 0
    variants = (('graph', True, None, False), ('eager', False, None, False))
    if fusable:
        variants += (('one_launch', True, None, True),)        # cb200_track_frame: the whole frame as one cooperative launch
    if fusable and full:
        variants += (('autograd_graph', True, False, False),)
    for name, use_graph, fused, persistent in variants:
        tr = FrameTracker(model, cfg, device=dev, use_graph=use_graph, fused=fused, persistent=persistent)
        for _ in range(3):
            tr.track(*draw(), c2w0)[0].cpu()
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        for _ in range(frames):
            tr.track(*draw(), c2w0)[0].cpu()          # the read-back is the per-frame sync
        out['ms_per_frame_' + name] = (time.perf_counter() - t0) * 1e3 / frames
        if name == 'graph':
            # the same loop with frame t+1's pixel draw + CPU gather done while the GPU tracks frame t (the draw does not depend
            # on the pose; same random.sample stream, same order)
            nxt = draw()
            torch.cuda.synchronize()
            t0 = time.perf_counter()
            for _ in range(frames):
                est = tr.track(*nxt, c2w0)[0]
                nxt = draw()
                est.cpu()
            out['ms_per_frame_graph_draw_overlapped'] = (time.perf_counter() - t0) * 1e3 / frames
    t0 = time.perf_counter()
    for _ in range(frames):
        draw()
    out['ms_per_frame_pixel_draw_and_gather'] = (time.perf_counter() - t0) * 1e3 / frames
    model.noise_mode, model._noise_counter = noise_state
    out['ms_per_frame'] = out['ms_per_frame_graph']
    return out


def time_mapping(model, cfg, dev, n_kf=8, calls=5):
    """Mapping ms/iteration (BASELINE.json configs 2-3: "mapping iters/sec"): global_BA (coslam.py:285-397) over a synthetic
    keyframe store of `n_kf` keyframes x `mapping.n_pixels` H W rays, `mapping.sample` + current-frame rays per iteration,
    `mapping.iters` iterations per call, through coslam_b200.mapping.BundleAdjuster (no autograd, one native call per
    iteration; host-side index draws bit-identical to the reference's `random.sample`).  Runs on a copy of the model."""
    import copy
    import random
    from coslam_b200 import optim as cb_optim
    from coslam_b200 import pose
    from coslam_b200.keyframe import KeyFrameDatabase
    from coslam_b200.mapping import BundleAdjuster
    model = copy.deepcopy(model)
    model.noise_mode, model._noise_counter = 'device', None
    m, cam = cfg['mapping'], cfg['cam']
    H, W = cam['H'], cam['W']
    R = int(H * W * m['n_pixels'])
    g = torch.Generator().manual_seed(9)
    bound = synth.bound_tensor(cfg)
    centre, ext = bound.mean(1), bound[:, 1] - bound[:, 0]
    dirs = synth.camera_ray_dirs(H, W, cam['fx'], cam['fy'], cam['cx'], cam['cy']).to(dev)
    db = KeyFrameDatabase(cfg, H, W, n_kf + 1, R, dev)
    poses = []
    random.seed(9)
    for k in range(n_kf + 1):
        r = torch.randn(1, 3, generator=g)
        poses.append(pose.pose_matrix(r / r.norm() * (0.2 + 0.02 * k), (centre + 0.1 * ext * (torch.rand(3, generator=g) - 0.5))[None])[0])
        batch = {'direction': dirs[None], 'rgb': torch.rand(1, H, W, 3, generator=g).to(dev),
                 'depth': (torch.rand(1, H, W, generator=g) * 3.0 + 0.5).to(dev), 'frame_id': torch.tensor([k * m['keyframe_every']])}
        if k < n_kf:
            db.add_keyframe(batch)
    current = torch.cat([batch['direction'], batch['rgb'], batch['depth'][..., None]], -1).reshape(-1, 7)
    kf_poses, cur_pose = torch.stack(poses[:n_kf]).to(dev), poses[n_kf].to(dev)
    opt = cb_optim.Adam([{'params': model.decoder.parameters(), 'weight_decay': 1e-6, 'lr': m['lr_decoder']},
                         {'params': model.embed_fn.parameters(), 'eps': 1e-15, 'lr': m['lr_embed']}], betas=(0.9, 0.99))
    ba = BundleAdjuster(model, cfg, db, opt, dev)
    for _ in range(2):
        ba.run(kf_poses, cur_pose, current)
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(calls):
        ba.run(kf_poses, cur_pose, current)[0].cpu()              # the updated poses are read back every call
    ms = (time.perf_counter() - t0) * 1e3 / (calls * m['iters'])
    n_rays = m['sample'] + max(m['sample'] // n_kf, m['min_pixels_cur'])
    return {'keyframes': n_kf, 'rays': n_rays, 'iters_per_call': m['iters'], 'ms_per_iter': ms, 'iters_per_s': 1e3 / ms,
            'smoothness_points': (cfg['training']['smooth_pts'] - 1) ** 3 if cfg['training']['smooth_weight'] > 0 else 0}


if __name__ == '__main__':
    a = parse()
    # stdout carries the ONE JSON line only: the model's construction messages (the reference prints the same ones,
    # model/scene_rep.py:36, encodings.py) and anything a library writes go to stderr
    sys.stdout.flush()
    _STDOUT = os.fdopen(os.dup(1), 'w')      # keep the real stdout for the JSON line ...
    os.dup2(2, 1)                            # ... and send file descriptor 1 to stderr too (NCCL prints its version banner there from C)
    sys.stdout = sys.stderr
    if a.impl == 'reference':
        run_reference(a)
    else:
        run_ours(a)
