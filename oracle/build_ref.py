"""TEST INFRASTRUCTURE ONLY - builds the UNMODIFIED reference extensions into oracle/_ref/.

Compiles the reference's own sources where they lie under /root/reference (never copied
into this repo) with torch.utils.cpp_extension for sm_100, producing

    oracle/_ref/roiaware_pool3d_cuda.so      (points_in_boxes_gpu / points_in_boxes_cpu)
    oracle/_ref/roipoint_pool3d_cuda.so      (forward)
    oracle/_ref/pointnet2_stack_cuda.so      (ball_query / group_points / FPS / voxel_query wrappers)
    oracle/_ref/pointnet2_batch_cuda.so      (batch-layout ball_query / group / gather / FPS / three_nn / interpolate)

    oracle/_ref/pywrap/*.pycode                 (the reference's own Python wrappers - pointnet2_utils.py, pointnet2_modules.py,
                                              roiaware_pool3d_utils.py, roipoint_pool3d_utils.py, box_utils.py, common_utils.py,
                                              voxel_query_utils.py, voxel_pool_modules.py - BYTE-COMPILED, unmodified: the
                                              drop-in tests import them over the stand-in extension modules)

oracle/_ref/ is git-ignored but travels to the GPU box with the gpurun snapshot.  Only
tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may
load these files; the product path never does.

Usage:  python oracle/build_ref.py [name ...]
"""
import glob
import os
import shutil
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
OUT = os.path.join(HERE, "_ref")
REF_OPS = "/root/reference/pcdet/ops"

EXTS = {
    "roiaware_pool3d_cuda": "roiaware_pool3d/src",
    "roipoint_pool3d_cuda": "roipoint_pool3d/src",
    "pointnet2_stack_cuda": "pointnet2/pointnet2_stack/src",
    "pointnet2_batch_cuda": "pointnet2/pointnet2_batch/src",
}


def built(name):
    return os.path.exists(os.path.join(OUT, name + ".so"))


def build_one(name):
    os.environ.setdefault("TORCH_CUDA_ARCH_LIST", "10.0")
    os.environ.setdefault("MAX_JOBS", "8")
    from torch.utils.cpp_extension import load

    src_dir = os.path.join(REF_OPS, EXTS[name])
    srcs = sorted(glob.glob(os.path.join(src_dir, "*.cpp")) + glob.glob(os.path.join(src_dir, "*.cu")))
    bdir = os.path.join(OUT, "_build_" + name)
    os.makedirs(bdir, exist_ok=True)
    load(name=name, sources=srcs, extra_include_paths=[src_dir], build_directory=bdir,
         extra_cflags=["-O2", "-w"], extra_cuda_cflags=["-O3", "-w"], verbose=False, is_python_module=False)
    shutil.copy(os.path.join(bdir, name + ".so"), os.path.join(OUT, name + ".so"))
    shutil.rmtree(bdir, ignore_errors=True)


# reference Python wrappers, byte-compiled as they are (module name inside the package -> source file)
PYWRAP = {
    "utils.common_utils": "utils/common_utils.py",
    "utils.box_utils": "utils/box_utils.py",
    "ops.roiaware_pool3d.roiaware_pool3d_utils": "ops/roiaware_pool3d/roiaware_pool3d_utils.py",
    "ops.roipoint_pool3d.roipoint_pool3d_utils": "ops/roipoint_pool3d/roipoint_pool3d_utils.py",
    "ops.pointnet2.pointnet2_stack.pointnet2_utils": "ops/pointnet2/pointnet2_stack/pointnet2_utils.py",
    "ops.pointnet2.pointnet2_stack.pointnet2_modules": "ops/pointnet2/pointnet2_stack/pointnet2_modules.py",
    "ops.pointnet2.pointnet2_stack.voxel_query_utils": "ops/pointnet2/pointnet2_stack/voxel_query_utils.py",
    "ops.pointnet2.pointnet2_stack.voxel_pool_modules": "ops/pointnet2/pointnet2_stack/voxel_pool_modules.py",
}


def build_pywrap():
    import py_compile
    out = os.path.join(OUT, "pywrap")
    os.makedirs(out, exist_ok=True)
    for mod, rel in PYWRAP.items():
        dst = os.path.join(out, mod + ".pycode")
        if not os.path.exists(dst):
            py_compile.compile(os.path.join(os.path.dirname(REF_OPS), rel), cfile=dst, doraise=True)


def main(names):
    if not os.path.isdir(REF_OPS):
        print("reference tree absent: using prebuilt oracle/_ref as-is")
        return 0
    os.makedirs(OUT, exist_ok=True)
    if not names:
        build_pywrap()
    for name in names or list(EXTS):
        if built(name):
            continue
        print("building reference extension", name, flush=True)
        build_one(name)
    return 0


if __name__ == "__main__":
    sys.exit(main(sys.argv[1:]))
