"""Loads the REFERENCE's own, unmodified Python wrappers (byte-compiled by oracle/build_ref.py into
oracle/_ref/pywrap/*.pycode (python bytecode; the extension is not .pyc because snapshots drop those); read straight from /root/reference when that tree is present) as a package whose extension
modules are chosen by the caller: the stand-ins of this repo (`which='new'`) or the reference's compiled extensions
(`which='ref'`).  Same Python, two native back ends - the drop-in contract of SURVEY 8b."""
import importlib.machinery
import importlib.util
import os
import sys
import types

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PYWRAP = os.path.join(ROOT, "oracle", "_ref", "pywrap")
REF = "/root/reference/pcdet"

MODULES = {   # module inside the package -> source path under /root/reference/pcdet
    "utils.common_utils": "utils/common_utils.py",
    "ops.roiaware_pool3d.roiaware_pool3d_utils": "ops/roiaware_pool3d/roiaware_pool3d_utils.py",
    "utils.box_utils": "utils/box_utils.py",                      # imports roiaware_pool3d_utils: after it
    "ops.roipoint_pool3d.roipoint_pool3d_utils": "ops/roipoint_pool3d/roipoint_pool3d_utils.py",
    "ops.pointnet2.pointnet2_stack.pointnet2_utils": "ops/pointnet2/pointnet2_stack/pointnet2_utils.py",
    "ops.pointnet2.pointnet2_stack.voxel_query_utils": "ops/pointnet2/pointnet2_stack/voxel_query_utils.py",
    "ops.pointnet2.pointnet2_stack.pointnet2_modules": "ops/pointnet2/pointnet2_stack/pointnet2_modules.py",
    "ops.pointnet2.pointnet2_stack.voxel_pool_modules": "ops/pointnet2/pointnet2_stack/voxel_pool_modules.py",
}
EXT = {   # extension module inside the package -> (stand-in module of this repo, reference .so name)
    "ops.roiaware_pool3d.roiaware_pool3d_cuda": ("multimodal_gar_b200.pcdet.ops.roiaware_pool3d.roiaware_pool3d_cuda", "roiaware_pool3d_cuda"),
    "ops.roipoint_pool3d.roipoint_pool3d_cuda": ("multimodal_gar_b200.pcdet.ops.roipoint_pool3d.roipoint_pool3d_cuda", "roipoint_pool3d_cuda"),
    "ops.pointnet2.pointnet2_stack.pointnet2_stack_cuda": ("multimodal_gar_b200.pcdet.ops.pointnet2.pointnet2_stack.pointnet2_stack_cuda", "pointnet2_stack_cuda"),
}


def available():
    return os.path.isdir(REF) or os.path.isdir(PYWRAP)


def _pkg(name):
    if name not in sys.modules:
        m = types.ModuleType(name)
        m.__path__ = []
        m.__package__ = name
        sys.modules[name] = m
        if "." in name:
            parent, leaf = name.rsplit(".", 1)
            setattr(_pkg(parent), leaf, m)
    return sys.modules[name]


def load(which):
    """-> dict: short name ('pointnet2_utils', ...) -> the reference module bound to the `which` extensions."""
    root = "refpcdet_" + which
    if root + ".ready" in sys.modules:
        return sys.modules[root + ".ready"].mods
    if "SharedArray" not in sys.modules:                  # common_utils.py:7 imports it at module level; never used here
        sys.modules["SharedArray"] = types.ModuleType("SharedArray")
    _pkg(root)
    for ext, (standin, so) in EXT.items():
        parent, leaf = (root + "." + ext).rsplit(".", 1)
        if which == "new":
            mod = importlib.import_module(standin)
        else:
            from oracle import oracle as O
            mod = O.load_ref(so)
            if mod is None:
                raise RuntimeError("oracle/_ref/%s.so is not built" % so)
        sys.modules[parent + "." + leaf] = mod
        setattr(_pkg(parent), leaf, mod)
    mods = {}
    for name, rel in MODULES.items():
        full = root + "." + name
        parent, leaf = full.rsplit(".", 1)
        _pkg(parent)
        src = os.path.join(REF, rel)
        if os.path.exists(src):
            loader = importlib.machinery.SourceFileLoader(full, src)
        else:
            loader = importlib.machinery.SourcelessFileLoader(full, os.path.join(PYWRAP, name + ".pycode"))
        spec = importlib.util.spec_from_loader(full, loader)
        m = importlib.util.module_from_spec(spec)
        m.__package__ = parent
        sys.modules[full] = m
        setattr(sys.modules[parent], leaf, m)
        import warnings
        with warnings.catch_warnings():
            warnings.simplefilter("ignore")               # the reference's docstrings carry invalid escape sequences
            loader.exec_module(m)
        mods[leaf] = m
    ready = types.ModuleType(root + ".ready")
    ready.mods = mods
    sys.modules[root + ".ready"] = ready
    return mods
