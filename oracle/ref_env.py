"""ORACLE tooling (test infrastructure) -- stand-ins for the third-party modules the reference imports at module level
but this image does not have, so that the reference's OWN files (utils.py, coslam.py, coslam_mp.py, mp_slam/*,
optimization/utils.py, tools/eval_ate.py) can be imported and run UNMODIFIED from /root/reference or baseline/_ref:

  pytorch3d.transforms   matrix_to_quaternion / quaternion_to_matrix / quaternion_to_axis_angle (optimization/utils.py:2):
                         restated from their published formulas in coslam_b200/pose.py (newer pytorch3d: non-negative real part)
  marching_cubes         marching_cubes(volume, isolevel, truncation=...) (utils.py:135): records the volume in
                         `LAST_VOLUME` and returns a one-triangle mesh -- the C++ extension itself is out of scope
  trimesh                Trimesh(vertices, faces, ...) with .export(path) (utils.py:176-199): a holder that writes a stub file
  matplotlib(.pyplot)    imported by utils.py:6 and tools/eval_ate.py, used only for figures: inert
  mathutils              tools/eval_ate.py:22, lazily, only for an unused branch: inert

Only tests/ and tools run under tests may import this file."""
import sys
import types

LAST_VOLUME = []           # volumes handed to marching_cubes.marching_cubes, newest last


class _Inert(object):
    """Accepts any attribute access / call chain (matplotlib figures)."""

    def __getattr__(self, name):
        if name.startswith('__'):
            raise AttributeError(name)
        return _Inert()

    def __call__(self, *a, **k):
        return _Inert()

    def __iter__(self):
        return iter(())


def _inert_module(name):
    m = types.ModuleType(name)

    def module_getattr(attr):
        if attr.startswith('__'):                          # inspect / importlib probe __file__, __path__, ...
            raise AttributeError(attr)
        return _Inert()
    m.__getattr__ = module_getattr
    return m


def install_stubs():
    import numpy as np
    from coslam_b200 import pose
    if 'pytorch3d' not in sys.modules or not hasattr(sys.modules.get('pytorch3d.transforms'), 'quaternion_to_axis_angle') \
            or sys.modules['pytorch3d.transforms'].matrix_to_quaternion is None:
        p3d, tr = types.ModuleType('pytorch3d'), types.ModuleType('pytorch3d.transforms')
        tr.quaternion_to_matrix = pose.quaternion_to_matrix
        tr.matrix_to_quaternion = pose.matrix_to_quaternion
        tr.quaternion_to_axis_angle = pose.quaternion_to_axis_angle
        tr.rotation_6d_to_matrix = None                       # imported, never called (optimization/utils.py)
        p3d.transforms = tr
        sys.modules['pytorch3d'], sys.modules['pytorch3d.transforms'] = p3d, tr
    if 'marching_cubes' not in sys.modules:
        mc = types.ModuleType('marching_cubes')

        def marching_cubes(volume, isolevel, truncation=None):
            LAST_VOLUME.append(np.array(volume, copy=True))
            # one triangle in lattice units, so that the caller's vertex pipeline (rescale, colour query) has something to chew on
            n = np.asarray(volume.shape[:3], dtype=np.float64)
            return np.stack([0.25 * n, 0.5 * n, 0.75 * n]), np.array([[0, 1, 2]], dtype=np.int64)
        mc.marching_cubes = marching_cubes
        sys.modules['marching_cubes'] = mc
    if 'trimesh' not in sys.modules:
        tm = types.ModuleType('trimesh')

        class Trimesh(object):
            def __init__(self, vertices=None, faces=None, process=False, vertex_colors=None, **kw):
                self.vertices, self.faces, self.vertex_colors = vertices, faces, vertex_colors
                self.vertex_normals = np.zeros((0, 3))

            def export(self, path):
                with open(path, 'w') as f:
                    f.write('stub mesh: %d vertices\n' % (0 if self.vertices is None else len(self.vertices)))
        tm.Trimesh = Trimesh
        sys.modules['trimesh'] = tm
    for name in ('matplotlib', 'matplotlib.pyplot', 'matplotlib.pylab', 'matplotlib.patches'):
        if name not in sys.modules:
            try:
                __import__(name)
            except ImportError:
                sys.modules[name] = _inert_module(name)
    if isinstance(sys.modules.get('matplotlib'), types.ModuleType) and not hasattr(sys.modules['matplotlib'], '__file__'):
        for sub in ('pyplot', 'pylab', 'patches'):
            setattr(sys.modules['matplotlib'], sub, sys.modules['matplotlib.' + sub])
    if 'mathutils' not in sys.modules:
        try:
            __import__('mathutils')
        except ImportError:
            mu = types.ModuleType('mathutils')

            class Matrix(object):
                """mathutils.Matrix(R).to_quaternion() as tools/eval_ate.py:22-26 uses it: (w, x, y, z) of a 3x3 rotation."""

                def __init__(self, rows):
                    self.m = np.asarray(rows, dtype=np.float64)

                def to_quaternion(self):
                    from scipy.spatial.transform import Rotation
                    x, y, z, w = Rotation.from_matrix(self.m).as_quat()
                    return np.array([w, x, y, z])
            mu.Matrix = Matrix
            sys.modules['mathutils'] = mu


def register_reference_packages(root):
    """The reference's top-level directories (`datasets`, `model`, `optimization`, `mp_slam`, `tools`) have no __init__.py, so a
    regular package of the same name anywhere on sys.path (HuggingFace `datasets` is installed in this image) would win the
    import.  Register them explicitly, as the packages they are in the reference's own environment."""
    import os
    for name in ('datasets', 'model', 'optimization', 'mp_slam', 'tools'):
        d = os.path.join(root, name)
        if os.path.isdir(d):
            pkg = types.ModuleType(name)
            pkg.__path__ = [d]
            sys.modules[name] = pkg
    if root not in sys.path:
        sys.path.insert(0, root)


def import_reference(name):
    """Imports module `name` of the reference tree (unmodified) with the stubs and the tinycudann shim registered."""
    import importlib
    from . import ref_loader, tcnn_shim
    install_stubs()
    sys.modules.setdefault('tinycudann', tcnn_shim)
    if ref_loader.REFERENCE_ROOT not in sys.path:
        sys.path.insert(0, ref_loader.REFERENCE_ROOT)
    return importlib.import_module(name)
