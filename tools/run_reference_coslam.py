"""Runs the reference's OWN, UNMODIFIED coslam.py (`CoSLAM.run()`, coslam.py:651-721, as `__main__`) on a synthetic Replica-format
sequence, with the scene model either

  --model ours        coslam_b200.scene_rep.JointEncoding   (the drop-in: INTEGRATION.md section A's one-line import swap, done
                      here without touching the file by registering the module `model.scene_rep` before coslam.py imports it)
  --model reference   the reference's model/scene_rep.py with oracle/tcnn_shim.py as `tinycudann` (pure torch on the same GPU)

Everything else is the reference's code from baseline/_ref (or /root/reference): coslam.py, config.py, datasets/dataset.py
(ReplicaDataset reads the frames this script writes: results/frame%06d.jpg, results/depth%06d.png, traj.txt), model/keyframe.py,
optimization/utils.py, utils.py (extract_mesh), tools/eval_ate.py.  Third-party modules this image lacks (pytorch3d, trimesh,
marching_cubes, matplotlib, mathutils) are the stand-ins of oracle/ref_env.py.  Test tooling: run by tests/test_reference_loop.py.

    python tools/run_reference_coslam.py --model ours --frames 30 --data /tmp/seq --out /tmp/run_ours.json

Prints / writes one JSON object: ATE RMSE (cm) of est_c2w_data against the generator's poses, the reference's own
pose_evaluation result, wall time, which JointEncoding class ran."""
import argparse
import json
import os
import runpy
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, 'tools'))

import numpy as np      # noqa: E402
import torch            # noqa: E402


def write_sequence(datadir, n_frames, scale, dev):
    """Sphere-traced frames of tools/slam_synth.py's room in the Replica layout the reference's loader expects
    (datasets/dataset.py:188-265): JPEG colour, 16-bit PNG depth (png_depth_scale 6553.5), traj.txt in the OpenCV convention
    (the loader flips the y and z columns back, :261-262)."""
    import cv2
    import slam_synth
    from coslam_b200 import synth
    cfg = slam_synth.scaled_config(synth.room0_config(), scale)
    cam = cfg['cam']
    for k in ('fx', 'fy', 'cx', 'cy'):                    # BaseDataset floors the intrinsics (datasets/dataset.py:50-53)
        cam[k] = float(cam[k] // 1)
    slam_synth.SHIFT.zero_()
    slam_synth.SCALE[0] = 1.0
    frames = slam_synth.make_frames(cfg, n_frames, dev)
    os.makedirs(os.path.join(datadir, 'results'), exist_ok=True)
    lines = []
    for f in frames:
        i = f['frame_id']
        rgb = (f['rgb'][0].clamp(0, 1) * 255.0).round().to(torch.uint8).cpu().numpy()
        depth = (f['depth'][0] * cam['png_depth_scale']).round().clamp(0, 65535).cpu().numpy().astype(np.uint16)
        cv2.imwrite(os.path.join(datadir, 'results', 'frame%06d.jpg' % i), cv2.cvtColor(rgb, cv2.COLOR_RGB2BGR), [cv2.IMWRITE_JPEG_QUALITY, 98])
        cv2.imwrite(os.path.join(datadir, 'results', 'depth%06d.png' % i), depth)
        c2w = f['c2w'].double().cpu().numpy().copy()
        c2w[:3, 1] *= -1
        c2w[:3, 2] *= -1
        lines.append(' '.join('%.17g' % v for v in c2w.reshape(-1)))
    with open(os.path.join(datadir, 'traj.txt'), 'w') as fh:
        fh.write('\n'.join(lines) + '\n')
    return cam


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument('--model', default='ours', choices=('ours', 'reference'))
    ap.add_argument('--frames', type=int, default=30)
    ap.add_argument('--scale', type=float, default=0.5, help='fraction of the 1200x680 Replica frame')
    ap.add_argument('--first-iters', type=int, default=200)
    ap.add_argument('--data', required=True, help='sequence directory (written if it has no traj.txt)')
    ap.add_argument('--workdir', required=True, help='output directory of the run')
    ap.add_argument('--out', default=None)
    a = ap.parse_args()
    from oracle import ref_env, ref_loader, tcnn_shim
    if not ref_loader.available():
        raise SystemExit('no reference tree: run `python -c "import __graft_entry__ as g; g.build()"` in the build container first')
    ref_root = ref_loader.REFERENCE_ROOT
    dev = 'cuda:0' if torch.cuda.is_available() else 'cpu'      # (the reference leg also runs on a CPU-only machine, slowly)
    if not os.path.isfile(os.path.join(a.data, 'traj.txt')):
        write_sequence(a.data, a.frames, a.scale, dev)
    from coslam_b200 import synth
    import slam_synth
    cam = slam_synth.scaled_config(synth.room0_config(), a.scale)['cam']

    # ---- the reference's environment: its packages, the third-party stand-ins, and the model under test ------------------
    ref_env.install_stubs()
    ref_env.register_reference_packages(ref_root)
    sys.modules['tinycudann'] = tcnn_shim
    if a.model == 'ours':
        import types
        from coslam_b200 import scene_rep
        swap = types.ModuleType('model.scene_rep')           # `from model.scene_rep import JointEncoding` -> the drop-in
        swap.JointEncoding = scene_rep.JointEncoding
        sys.modules['model.scene_rep'] = swap
    os.makedirs(a.workdir, exist_ok=True)
    cfg_path = os.path.join(a.workdir, 'synthetic_room0.yaml')
    with open(cfg_path, 'w') as fh:                          # a user's scene YAML on top of the shipped Replica one
        fh.write('inherit_from: configs/Replica/room0.yaml\n'
                 'data:\n  datadir: %s\n  output: %s\n  exp_name: run\n  num_workers: 0\n'
                 'cam:\n  H: %d\n  W: %d\n  fx: %r\n  fy: %r\n  cx: %r\n  cy: %r\n'
                 'tracking:\n  ignore_edge_W: %d\n  ignore_edge_H: %d\n'
                 'mapping:\n  first_mesh: False\n  first_iters: %d\n'
                 'mesh:\n  voxel_final: 0.1\n'
                 % (a.data, os.path.join(a.workdir, 'output'), cam['H'], cam['W'], cam['fx'], cam['fy'], cam['cx'], cam['cy'],
                    int(20 * a.scale), int(20 * a.scale), a.first_iters))
    # coslam.py writes its checkpoint / pose files into <output>/<exp_name>, which it only creates as a side effect of the first
    # mesh (extract_mesh's makedirs); with `first_mesh: False` the user has to provide it
    os.makedirs(os.path.join(a.workdir, 'output', 'run'), exist_ok=True)
    os.chdir(ref_root)                                       # `inherit_from` paths are relative to the reference root
    sys.argv = ['coslam.py', '--config', cfg_path]
    sync = torch.cuda.synchronize if torch.cuda.is_available() else (lambda: None)
    sync()
    t0 = time.perf_counter()
    g = runpy.run_path(os.path.join(ref_root, 'coslam.py'), run_name='__main__')      # UNMODIFIED: argparse, CoSLAM(cfg), seed_everything(0), run()
    sync()
    wall = time.perf_counter() - t0
    slam = g['slam']
    n = len(slam.est_c2w_data)
    est = torch.stack([slam.est_c2w_data[i].detach().cpu() for i in range(n)])
    gt = torch.stack([slam.pose_gt[i].cpu() for i in range(n)])
    ate = float(torch.sqrt(((est[:, :3, 3] - gt[:, :3, 3]) ** 2).sum(-1).mean()))
    out_dir = os.path.join(a.workdir, 'output', 'run')
    ref_eval = None
    try:                                                     # what the reference's own pose_evaluation wrote (Horn-aligned ATE)
        last = open(os.path.join(out_dir, 'output.txt')).read().strip().splitlines()[-1]
        import numpy  # noqa: F401  (the line holds `np.float64(...)` reprs under numpy 2)
        ref_eval = float(eval(last, {'np': np, 'numpy': np, 'array': np.array, 'float64': np.float64})['absolute_translational_error.rmse'])
    except Exception as e:
        ref_eval = 'unreadable: %s' % e
    model = slam.model
    res = {'model': a.model, 'model_class': type(model).__module__ + '.' + type(model).__name__, 'frames': n,
           'frame': '%dx%d' % (cam['W'], cam['H']), 'ate_rmse_cm': 100 * ate,
           'reference_pose_evaluation_ate_rmse_cm': 100 * ref_eval if isinstance(ref_eval, float) else ref_eval,
           'wall_s': wall, 'frames_per_s': (n - 1) / wall, 'keyframes': len(slam.keyframeDatabase.frame_ids),
           'checkpoint_written': os.path.isfile(os.path.join(out_dir, 'checkpoint%d.pt' % (n - 1))),
           'mesh_volume_queried': len(ref_env.LAST_VOLUME) > 0,
           'embed_params_class': type(model.embed_fn).__module__}
    print(json.dumps(res))
    if a.out:
        with open(a.out, 'w') as fh:
            json.dump(res, fh)


if __name__ == '__main__':
    main()
