"""ORACLE tooling (test infrastructure) -- imports the UNMODIFIED reference model from /root/reference (build container) or
from baseline/_ref (the unmodified copy `__graft_entry__.build()` installs there; git-ignored, shipped to the GPU box by
gpurun) with oracle/tcnn_shim.py registered as `tinycudann`.  Used by oracle/gen_golden.py, the CPU cross-check tests
(skipped when no tree is present), bench.py's CPU arm (`--impl reference`, `cpu_baseline`: kind "reference") and the
coslam.py drop-in test.  /root/reference itself is never read on the GPU box."""
import os
import sys

_HERE = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _default_root():
    # /root/reference in the build container; baseline/_ref (the unmodified copy __graft_entry__.build() installs, which
    # travels with gpurun) on the GPU box
    if os.path.isfile('/root/reference/model/scene_rep.py'):
        return '/root/reference'
    return os.path.join(_HERE, 'baseline', '_ref')


REFERENCE_ROOT = os.environ.get('COSLAM_REFERENCE_ROOT') or _default_root()


def available():
    return os.path.isfile(os.path.join(REFERENCE_ROOT, 'model', 'scene_rep.py'))


def load_reference_model_class():
    """Returns the reference's JointEncoding class (model/scene_rep.py:10)."""
    if not available():
        raise RuntimeError('reference tree not present at %s' % REFERENCE_ROOT)
    from . import tcnn_shim
    sys.modules['tinycudann'] = tcnn_shim
    if REFERENCE_ROOT not in sys.path:
        sys.path.insert(0, REFERENCE_ROOT)
    from model.scene_rep import JointEncoding      # noqa: the reference's own, unmodified
    return JointEncoding


def load_reference_config(rel_path):
    """config.py:4-33 load_config, run from the reference root because inherit_from paths are relative."""
    if REFERENCE_ROOT not in sys.path:
        sys.path.insert(0, REFERENCE_ROOT)
    import config as ref_config
    cwd = os.getcwd()
    os.chdir(REFERENCE_ROOT)
    try:
        return ref_config.load_config(rel_path)
    finally:
        os.chdir(cwd)
