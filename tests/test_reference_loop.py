"""The reference's own, UNMODIFIED coslam.py (`CoSLAM.run()`, coslam.py:651-721, executed as `__main__`) on top of the drop-in
JointEncoding -- SURVEY.md section 8(b): "keeps coslam.py ... unchanged" as a fact, not a claim.

tools/run_reference_coslam.py writes a synthetic Replica-format sequence (loaded by the reference's ReplicaDataset), then runs
baseline/_ref/coslam.py twice in fresh processes on the same files: once with `model.scene_rep` resolved to
coslam_b200.scene_rep (INTEGRATION.md section A's one-line swap), once with the reference's own scene_rep.py on the pure-torch
tinycudann stand-in.  Both must finish (first-frame mapping, tracking every frame, global BA every 5th, keyframes, checkpoint,
final mesh query, the reference's pose_evaluation) and end with the same trajectory quality: north_star's ATE gate, 0.2 cm."""
import json
import os
import subprocess
import sys

import pytest

from oracle import ref_loader

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
pytestmark = [pytest.mark.gpu,
              pytest.mark.skipif(not ref_loader.available(), reason='baseline/_ref not installed (__graft_entry__.build() in the build container)')]


def _run(model, data, workdir, frames=30):
    out = os.path.join(workdir, 'result.json')
    cmd = [sys.executable, os.path.join(ROOT, 'tools', 'run_reference_coslam.py'), '--model', model, '--frames', str(frames),
           '--scale', '0.5', '--data', data, '--workdir', workdir, '--out', out]
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=1500)
    assert r.returncode == 0, 'coslam.py (%s) failed:\n%s\n%s' % (model, r.stdout[-3000:], r.stderr[-3000:])
    return json.load(open(out))


def test_unmodified_coslam_py_runs_on_the_drop_in(tmp_path):
    data = str(tmp_path / 'room0_synth')
    ours = _run('ours', data, str(tmp_path / 'ours'))
    ref = _run('reference', data, str(tmp_path / 'reference'))
    os.makedirs(os.path.join(ROOT, 'gpurun_out'), exist_ok=True)
    with open(os.path.join(ROOT, 'gpurun_out', 'reference_coslam_loop.json'), 'w') as f:
        json.dump({'ours': ours, 'reference_model': ref}, f, indent=1)
    assert ours['model_class'] == 'coslam_b200.scene_rep.JointEncoding' and ours['embed_params_class'].startswith('coslam_b200')
    assert ref['model_class'] == 'model.scene_rep.JointEncoding'
    for r in (ours, ref):
        assert r['frames'] == 30 and r['keyframes'] == 6 and r['checkpoint_written'] and r['mesh_volume_queried']
        assert r['ate_rmse_cm'] < 2.0, r                       # both track the 30-frame sequence (static-pose ATE is ~15 cm)
    assert abs(ours['ate_rmse_cm'] - ref['ate_rmse_cm']) < 0.2, (ours, ref)
    assert abs(ours['reference_pose_evaluation_ate_rmse_cm'] - ref['reference_pose_evaluation_ate_rmse_cm']) < 0.2, (ours, ref)


def _run_mp(model, data, workdir, frames=30, timeout=600):
    import signal
    out = os.path.join(workdir, 'result.json')
    os.makedirs(workdir, exist_ok=True)
    cmd = [sys.executable, os.path.join(ROOT, 'tools', 'run_reference_coslam_mp.py'), '--model', model, '--frames', str(frames),
           '--scale', '0.5', '--data', data, '--workdir', workdir, '--out', out]
    log = open(os.path.join(workdir, 'log.txt'), 'w')
    p = subprocess.Popen(cmd, stdout=log, stderr=subprocess.STDOUT, start_new_session=True)      # own process group: the workers die with it
    try:
        rc = p.wait(timeout=timeout)
    except subprocess.TimeoutExpired:
        os.killpg(p.pid, signal.SIGKILL)
        p.wait()
        rc = 'timeout after %d s' % timeout
    finally:
        try:
            os.killpg(p.pid, signal.SIGKILL)                  # a worker that outlived the parent (one side crashed, the other spins)
        except ProcessLookupError:
            pass
    log.close()
    tail = open(os.path.join(workdir, 'log.txt')).read()[-4000:]
    assert rc == 0, 'coslam_mp.py (%s) failed (%s):\n%s' % (model, rc, tail)
    return json.load(open(out))


def test_unmodified_coslam_mp_py_runs_on_the_drop_in(tmp_path):
    """coslam_mp.py:295-320 unchanged: mapper process + tracker process (spawn) on ONE GPU, sharing the drop-in model through
    `.share_memory()` / pickling / `copy.deepcopy` (mp_slam/tracker.py:27), the pose tensors and the two progress counters."""
    data = str(tmp_path / 'room0_synth')
    ours = _run_mp('ours', data, str(tmp_path / 'mp_ours'))
    os.makedirs(os.path.join(ROOT, 'gpurun_out'), exist_ok=True)
    with open(os.path.join(ROOT, 'gpurun_out', 'reference_coslam_mp_loop.json'), 'w') as f:
        json.dump({'ours': ours}, f, indent=1)
    assert ours['model_class'] == 'coslam_b200.scene_rep.JointEncoding'
    assert ours['tracked_to'] == 29 and ours['all_poses_filled'], ours
    assert ours['ate_rmse_cm'] < 2.0, ours
