"""Run-to-run spread of the synthetic SLAM loop: two CUDA-path runs with identical seeds (atomics order differs), the oracle
arm and a device-noise run.  `python tests/traj_probe.py <frames> <first_iters> <scale>`; numbers quoted in DESIGN.md section 5."""
import sys, torch
sys.path.insert(0,'tests'); sys.path.insert(0,'tools'); sys.path.insert(0,'.')
import test_slam_gpu as T
slam_synth, OracleSLAM = T._oracle_slam_class()
from coslam_b200 import synth
dev='cuda:0'
n_frames, first_iters, scale = int(sys.argv[1]), int(sys.argv[2]), float(sys.argv[3])
cfg = lambda: slam_synth.scaled_config(synth.room0_config(), scale)
frames = slam_synth.make_frames(cfg(), n_frames, dev)
gt = torch.stack([f['c2w'] for f in frames]).cpu()
res = {}
w0 = None
for name in ('ours_a', 'ours_b', 'oracle', 'ours_dev'):
    if name == 'oracle':
        s = OracleSLAM(cfg(), dev, n_frames)
        s.model.load_weights(*w0)
    else:
        s = slam_synth.SynthSLAM(cfg(), dev, n_frames, noise='device' if name == 'ours_dev' else 'host')
        if w0 is None:
            w0 = [p.detach().clone() for p in s.model._fused()]
        else:
            with torch.no_grad():
                for p, q in zip(s.model._fused(), w0):
                    p.copy_(q)
    slam_synth.seed_everything(0)
    res[name] = s.run([dict(f) for f in frames], first_iters).cpu()
    print(name, 'ATE cm %.3f' % (100 * slam_synth.ate_rmse(res[name], gt)), flush=True)
for a in res:
    for b in res:
        if a < b:
            print(a, b, 'between cm %.3f' % (100 * slam_synth.ate_rmse(res[a], res[b])))
print('drift cm %.2f' % (100 * slam_synth.ate_rmse(gt[:1].expand_as(gt), gt)))
