"""ORACLE tooling (test infrastructure) -- imports the UNMODIFIED reference model from /root/reference with
oracle/tcnn_shim.py registered as `tinycudann`.  Works only in the build container: /root/reference does not
exist on the GPU box, so nothing under tests/ (-m gpu), smoke() or bench.py may call this at run time.  It is
used by oracle/gen_golden.py and by the CPU-only cross-check tests, which skip when the tree is absent."""
import os
import sys

REFERENCE_ROOT = os.environ.get('COSLAM_REFERENCE_ROOT', '/root/reference')


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
