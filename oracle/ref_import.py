"""ORACLE support -- import the real reference (read-only) in the build container.

Used only by ``tests/golden/make_golden.py`` (fixture generation) and by the
``-m "not gpu"`` tests that check the oracle against a live run of the reference
when ``/root/reference`` (or ``$VFDEPTH_REF``) is mounted.  It does not exist on
the GPU box, so nothing on a GPU path may call this.

The reference imports three packages that are not installed here at module top
(SURVEY.md 8c): ``pytorch3d.transforms`` (volumetric_fusionnet.py:5,
geometry_util.py:5, multi_cam_loss.py:3), ``tensorboardX`` and
``matplotlib.pyplot`` (utils/logger.py, utils/visualize.py).  They are stubbed in
``sys.modules``; ``models/__init__.py`` (which drags in the dataset stack) is
bypassed by registering bare package objects that point at the real directories.
Nothing is copied and the reference tree is never written to.
"""
import importlib
import os
import sys
import types

REF_ROOT = os.environ.get('VFDEPTH_REF', '/root/reference')


def available():
    return os.path.isfile(os.path.join(REF_ROOT, 'network', 'volumetric_fusionnet.py'))


def _stub(name, **attrs):
    mod = types.ModuleType(name)
    for k, v in attrs.items():
        setattr(mod, k, v)
    sys.modules[name] = mod
    return mod


def _bare_package(name, path):
    pkg = types.ModuleType(name)
    pkg.__path__ = [path]
    sys.modules[name] = pkg
    return pkg


_loaded = {}


def load():
    """-> namespace with the reference's hot-path classes/functions."""
    if _loaded:
        return types.SimpleNamespace(**_loaded)
    if not available():
        raise RuntimeError('reference tree not mounted at %s' % REF_ROOT)
    from vfdepth_b200.rig import axis_angle_to_matrix

    def _unused(*a, **k):
        raise NotImplementedError('stubbed third-party symbol')

    if 'pytorch3d' not in sys.modules:
        _stub('pytorch3d')
        _stub('pytorch3d.transforms', axis_angle_to_matrix=axis_angle_to_matrix, matrix_to_euler_angles=_unused)
    if 'tensorboardX' not in sys.modules:
        _stub('tensorboardX', SummaryWriter=object)
    try:
        import matplotlib.pyplot  # noqa: F401
    except Exception:
        _stub('matplotlib')
        _stub('matplotlib.pyplot')
    if REF_ROOT not in sys.path:
        sys.path.insert(0, REF_ROOT)
    # models/__init__.py imports vfdepth.py -> dataset -> packnet-sfm; bypass it.
    for pkg in ('models', 'models.geometry', 'models.losses'):
        _bare_package(pkg, os.path.join(REF_ROOT, *pkg.split('.')))
    cwd = os.getcwd()
    os.chdir(REF_ROOT)           # utils/__init__.py appends relative lib paths
    try:
        vf = importlib.import_module('network.volumetric_fusionnet')
        geo = importlib.import_module('models.geometry.geometry_util')
        vr = importlib.import_module('models.geometry.view_rendering')
        pose = importlib.import_module('models.geometry.pose')
        lu = importlib.import_module('models.losses.loss_util')
        mcl = importlib.import_module('models.losses.multi_cam_loss')
    finally:
        os.chdir(cwd)
    _loaded.update(VFNet=vf.VFNet, Projection=geo.Projection, ViewRendering=vr.ViewRendering, Pose=pose.Pose,
                   loss_util=lu, MultiCamLoss=mcl.MultiCamLoss)
    return types.SimpleNamespace(**_loaded)


def load_algo():
    """-> the reference's ``VFDepthAlgo`` class (models/vfdepth.py), for its stateless helper methods such as
    ``to_depth`` (:277-288).  ``dataset`` (packnet-sfm stack, absent) is stubbed; the class is never constructed --
    methods are called unbound on a ``types.SimpleNamespace`` carrying the config attributes they read."""
    if 'VFDepthAlgo' in _loaded:
        return _loaded['VFDepthAlgo']
    ref = load()

    def _unused(*a, **k):
        raise NotImplementedError('stubbed third-party symbol')

    if 'dataset' not in sys.modules:
        _stub('dataset', construct_dataset=_unused)
    cwd = os.getcwd()
    os.chdir(REF_ROOT)
    try:
        importlib.import_module('network')
        geo_pkg, loss_pkg = sys.modules['models.geometry'], sys.modules['models.losses']
        geo_pkg.Pose, geo_pkg.ViewRendering = ref.Pose, ref.ViewRendering
        loss_pkg.MultiCamLoss = ref.MultiCamLoss
        loss_pkg.SingleCamLoss = importlib.import_module('models.losses.single_cam_loss').SingleCamLoss
        loss_pkg.DepthSynLoss = importlib.import_module('models.losses.depth_synthesis_loss').DepthSynLoss
        algo = importlib.import_module('models.vfdepth').VFDepthAlgo
    finally:
        os.chdir(cwd)
    _loaded['VFDepthAlgo'] = algo
    return algo
