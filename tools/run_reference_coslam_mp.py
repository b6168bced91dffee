"""Runs the reference's OWN, UNMODIFIED coslam_mp.py (tracker process + mapper process, coslam_mp.py:295-320, mp_slam/tracker.py,
mp_slam/mapper.py) on the synthetic Replica-format sequence of tools/run_reference_coslam.py, with `--model ours` (the drop-in
JointEncoding) or `--model reference` (the reference model on the pure-torch tinycudann stand-in).

coslam_mp.py starts its two workers with the `spawn` method and pickles the bound methods `CoSLAM.tracking` / `CoSLAM.mapping`:
each child re-imports the parent's main module and looks `CoSLAM` up in it.  So this launcher IS the main module: at import
time (in the parent and again in each spawned child, which inherit the environment variables set below) it prepares the
reference's environment -- its package directories, the third-party stand-ins of oracle/ref_env.py, and the model swap of
INTEGRATION.md section A -- and then executes the source text of coslam_mp.py, as it is on disk, in this module's namespace.  In
the parent `__name__ == '__main__'`, so the file's own main block runs (argparse, config, CoSLAM(cfg), slam.run()); in the
children it only defines the classes.

    python tools/run_reference_coslam_mp.py --model ours --frames 30 --data /tmp/seq --workdir /tmp/mp_ours --out /tmp/mp_ours.json
"""
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for _p in (ROOT, os.path.join(ROOT, 'tools')):
    if _p not in sys.path:
        sys.path.insert(0, _p)


def _prepare_reference_environment(model_kind):
    import types
    from oracle import ref_env, ref_loader, tcnn_shim
    ref_env.install_stubs()
    ref_env.register_reference_packages(ref_loader.REFERENCE_ROOT)
    sys.modules['tinycudann'] = tcnn_shim
    if model_kind == 'ours':
        from coslam_b200 import scene_rep
        swap = types.ModuleType('model.scene_rep')
        swap.JointEncoding = scene_rep.JointEncoding
        sys.modules['model.scene_rep'] = swap
    return ref_loader.REFERENCE_ROOT


def _parent_setup():
    import argparse
    ap = argparse.ArgumentParser()
    ap.add_argument('--model', default='ours', choices=('ours', 'reference'))
    ap.add_argument('--frames', type=int, default=30)
    ap.add_argument('--scale', type=float, default=0.5)
    ap.add_argument('--first-iters', type=int, default=200)
    ap.add_argument('--data', required=True)
    ap.add_argument('--workdir', required=True)
    ap.add_argument('--out', default=None)
    a = ap.parse_args()
    import torch
    import run_reference_coslam as single
    import slam_synth
    from coslam_b200 import synth
    from oracle import ref_loader
    if not ref_loader.available():
        raise SystemExit('no reference tree (baseline/_ref)')
    if not os.path.isfile(os.path.join(a.data, 'traj.txt')):
        single.write_sequence(a.data, a.frames, a.scale, 'cuda:0' if torch.cuda.is_available() else 'cpu')
    cam = slam_synth.scaled_config(synth.room0_config(), a.scale)['cam']
    os.makedirs(os.path.join(a.workdir, 'output', 'run'), exist_ok=True)
    cfg_path = os.path.join(a.workdir, 'synthetic_room0.yaml')
    with open(cfg_path, 'w') as fh:
        fh.write('inherit_from: configs/Replica/room0.yaml\n'
                 'data:\n  datadir: %s\n  output: %s\n  exp_name: run\n  num_workers: 0\n'
                 'cam:\n  H: %d\n  W: %d\n  fx: %r\n  fy: %r\n  cx: %r\n  cy: %r\n'
                 'tracking:\n  ignore_edge_W: %d\n  ignore_edge_H: %d\n'
                 'mapping:\n  first_mesh: False\n  first_iters: %d\n'
                 'mesh:\n  voxel_final: 0.1\n'
                 % (a.data, os.path.join(a.workdir, 'output'), cam['H'], cam['W'], cam['fx'], cam['fy'], cam['cx'], cam['cy'],
                    int(20 * a.scale), int(20 * a.scale), a.first_iters))
    os.environ['CB200_REF_MP_MODEL'] = a.model
    sys.argv = ['coslam_mp.py', '--config', cfg_path]
    return a, cam


_ARGS = None
if __name__ == '__main__':
    _ARGS, _CAM = _parent_setup()

if os.environ.get('CB200_REF_MP_MODEL'):
    _REF_ROOT = _prepare_reference_environment(os.environ['CB200_REF_MP_MODEL'])
    if __name__ == '__main__':
        os.chdir(_REF_ROOT)                                   # `inherit_from` and shutil.copy("coslam.py", ...) are relative
    _T0 = time.perf_counter()
    _path = os.path.join(_REF_ROOT, 'coslam_mp.py')
    exec(compile(open(_path).read(), _path, 'exec'), globals())      # the file as it is on disk; its main block runs in the parent only

if __name__ == '__main__':
    import numpy as np
    import torch as _torch
    wall = time.perf_counter() - _T0
    a = _ARGS
    s = globals()['slam']
    n = a.frames
    est = s.est_c2w_data.detach().cpu()[:n]
    gt = s.pose_gt[:n]
    ate = float(_torch.sqrt(((est[:, :3, 3] - gt[:, :3, 3]) ** 2).sum(-1).mean()))
    out_dir = os.path.join(a.workdir, 'output', 'run')
    try:
        last = open(os.path.join(out_dir, 'output.txt')).read().strip().splitlines()[-1]
        ref_eval = 100 * float(eval(last, {'np': np, 'numpy': np, 'array': np.array, 'float64': np.float64})['absolute_translational_error.rmse'])
    except Exception as e:
        ref_eval = 'unreadable: %s' % e
    res = {'script': 'coslam_mp.py', 'model': a.model, 'model_class': type(s.model).__module__ + '.' + type(s.model).__name__,
           'frames': n, 'frame': '%dx%d' % (_CAM['W'], _CAM['H']), 'ate_rmse_cm': 100 * ate,
           'reference_pose_evaluation_ate_rmse_cm': ref_eval, 'wall_s': wall, 'frames_per_s': (n - 1) / wall,
           'tracked_to': int(s.tracking_idx[0]), 'mapped_to': int(s.mapping_idx[0]),
           'all_poses_filled': bool((est[:, 3, 3] == 1).all())}
    print(json.dumps(res))
    if a.out:
        with open(a.out, 'w') as fh:
            json.dump(res, fh)
