"""Generates tests/golden/ref_gpu.npz from the UNMODIFIED reference CUDA extensions (oracle/_ref,
built by oracle/build_ref.py from /root/reference sources) on a GPU box:

    gpurun -- 'python tests/golden/make_ref_gpu_golden.py gpurun_out/ref_gpu.npz'

then copy gpurun_out/ref_gpu.npz to tests/golden/.  The file pins the CPU oracle (and through it
the CUDA path) to the reference's own outputs; CPU tests read it, nothing reads /root/reference.
"""
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import cases  # noqa: E402
from oracle import oracle as O  # noqa: E402


def main(out_path):
    dev = torch.device("cuda:0")
    T = lambda a: torch.from_numpy(np.ascontiguousarray(a)).to(dev)
    roiaware = O.load_ref("roiaware_pool3d_cuda")
    roipoint = O.load_ref("roipoint_pool3d_cuda")
    pn2 = O.load_ref("pointnet2_stack_cuda")
    g = {"torch": np.array(torch.__version__), "cuda": np.array(torch.version.cuda),
         "gpu": np.array(torch.cuda.get_device_name(0))}

    xyz, feat, boxes = cases.frame_case()
    B, N, M = xyz.shape[0], xyz.shape[1], boxes.shape[1]
    g.update(f_xyz=xyz, f_feat=feat, f_boxes=boxes)
    out = torch.full((B, N), -1, dtype=torch.int32, device=dev)
    roiaware.points_in_boxes_gpu(T(boxes), T(xyz), out)
    g["pib_gpu"] = out.cpu().numpy()
    m = torch.zeros((M, N), dtype=torch.int32)
    roiaware.points_in_boxes_cpu(torch.from_numpy(boxes[0]), torch.from_numpy(xyz[0]), m)
    g["pib_cpu_frame0"] = m.numpy()
    for tag, extra, S in (("e0", (0.0, 0.0, 0.0), 128), ("e2", (0.2, 0.2, 0.2), 64)):
        big = O.enlarge_box3d(boxes, extra)
        pooled = torch.zeros((B, M, S, 4), device=dev)
        empty = torch.zeros((B, M), dtype=torch.int32, device=dev)
        roipoint.forward(T(xyz), T(big), T(feat), pooled, empty)
        g["pool_%s" % tag] = pooled.cpu().numpy()
        g["pool_empty_%s" % tag] = empty.cpu().numpy()

    for n, m_s in ((128, 64), (200, 64), (512, 512), (1500, 96), (2500, 40)):
        c = cases.cloud_with_ties(n, 3, n)
        temp = torch.full((3, n), 1e10, device=dev)
        idx = torch.zeros((3, m_s), dtype=torch.int32, device=dev)
        pn2.farthest_point_sampling_wrapper(3, n, m_s, T(c), temp, idx)
        g["fps_%d" % n] = idx.cpu().numpy()
        g["fps_temp_%d" % n] = temp.cpu().numpy()

    rx, rcnt, rnew, rncnt, rfeat = cases.ragged_case()
    g.update(r_xyz=rx, r_cnt=rcnt, r_new=rnew, r_ncnt=rncnt, r_feat=rfeat)
    pos = rcnt > 0
    npoint = np.array([32, 64, 16, 40], dtype=np.int32)
    temp = torch.full((rx.shape[0],), 1e10, device=dev)
    idx = torch.zeros((int(npoint.sum()),), dtype=torch.int32, device=dev)
    pn2.stack_farthest_point_sampling_wrapper(T(rx), temp, T(rcnt[pos]), idx, T(npoint))
    g["sfps_npoint"] = npoint
    g["sfps"] = idx.cpu().numpy()
    for r, K in ((0.3, 16), (0.6, 32), (0.05, 8)):
        bidx = torch.zeros((rnew.shape[0], K), dtype=torch.int32, device=dev)
        pn2.ball_query_wrapper(len(rcnt), rnew.shape[0], r, K, T(rnew), T(rncnt), T(rx), T(rcnt), bidx)
        g["bq_%g_%d" % (r, K)] = bidx.cpu().numpy()
        fixed = bidx.clone()
        fixed[fixed[:, 0] == -1] = 0
        outg = torch.zeros((rnew.shape[0], 2, K), device=dev)
        pn2.group_points_wrapper(len(rcnt), rnew.shape[0], 2, K, T(rfeat), T(rcnt), fixed, T(rncnt), outg)
        g["grp_%g_%d" % (r, K)] = outg.cpu().numpy()
    torch.cuda.synchronize()
    np.savez_compressed(out_path, **g)
    print("wrote", out_path, os.path.getsize(out_path), "bytes")


if __name__ == "__main__":
    main(sys.argv[1] if len(sys.argv) > 1 else os.path.join(ROOT, "tests", "golden", "ref_gpu.npz"))
