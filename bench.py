"""bench.py - JRDB-size LiDAR frames/s through box-pool + FPS + ball query + group + MLP on B200.

    python bench.py [--gpus N --steps K --warmup W]             (native arm, N=1 by default)
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N ... bench.py --gpus N ...
    python bench.py --impl reference ...                         (the path on the host cores)

A step = one pass of the hot path over `frames_per_step` synthetic JRDB-shaped frames of
BASELINE.json configs[1] (65 536 points, 40 person boxes, S=512, P=512, r=0.3, K=32, MLP 4-64-128).
Rank 0 prints ONE JSON line.  Frames are independent: every rank runs its own frames (weak scaling),
no collective in the data path; timing is CUDA events, max over ranks.
"""
import argparse
import json
import os
import statistics
import subprocess
import sys
import tempfile
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

CFG = dict(N=65536, M=40, S=512, P=512, radius=0.3, K=32, C_in=1, mlp=(64, 128))
METRIC = "JRDB-size LiDAR frames/s (box-pool+FPS+ballquery+group+MLP)"


def algorithmic_bytes_per_frame(c=CFG, keep_sa=True):
    """compulsory traffic, every tensor touched once (fp32/int32).  keep_sa=False: the
    (M*P, C2) per-centroid SA features stay on chip (max over a person's centroids inside the MLP kernel) and only
    the (M, C2) person features are written."""
    N, M, S, P, K, C2 = c["N"], c["M"], c["S"], c["P"], c["K"], c["mlp"][-1]
    return 16 * N + 28 * M + 4 * N + 16 * M * S + 4 * M * P + 4 * M * P * K + 4 * M * C2 * (P if keep_sa else 1)


def mlp_flops_per_frame(c=CFG):
    ch = [3 + c["C_in"]] + list(c["mlp"])
    return 2 * c["M"] * c["P"] * c["K"] * sum(a * b for a, b in zip(ch, ch[1:]))


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="native", choices=["native", "reference"])
    ap.add_argument("--frames-per-step", type=int, default=64)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--cpu-frames", type=int, default=0, help="frames in the cpu_baseline sample (0: auto)")
    return ap.parse_args()


def peaks():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(path):
        p = json.load(open(path))
        return dict(hbm=float(p["hbm_gbs"]), bf16=float(p["bf16_tflops"]), bf16_sustained=float(
            p.get("bf16_tflops_sustained", p["bf16_tflops"])), source="measured (MEASURED_PEAKS.json)")
    return dict(hbm=6650.0, bf16=1590.0, bf16_sustained=1400.0, source="fallback (B200_PROFILING.md)")


def ncu_traffic(stage, F):
    """dram__bytes_read.sum + dram__bytes_write.sum of the stage's kernel for one launch, from the committed
    `ncu --set full` capture of this same command (64 frames per launch); None when no capture matches."""
    path = os.path.join(ROOT, "profiles", "r01_traffic.json")
    try:
        t = json.load(open(path))
        return t["bytes_per_launch"][stage] if t.get("frames_per_step") == F else None
    except (OSError, KeyError, ValueError):
        return None


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled every 200 ms while the timed region runs."""
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.f = tempfile.NamedTemporaryFile("w+", suffix=".csv", delete=False)
        self.p = None
        try:
            self.p = subprocess.Popen(["nvidia-smi", "-i", str(index), "--query-gpu=" + self.Q,
                                       "--format=csv,noheader,nounits", "-lms", "200"], stdout=self.f,
                                      stderr=subprocess.DEVNULL)
        except OSError:
            pass

    def stop(self):
        out = dict(sm_mhz=None, sm_max_mhz=None, reasons=[], samples=0)
        if self.p is None:
            return out
        time.sleep(0.25)
        self.p.terminate()
        try:
            self.p.wait(timeout=5)
        except subprocess.TimeoutExpired:
            self.p.kill()
        self.f.seek(0)
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for line in self.f.read().splitlines():
            parts = [x.strip() for x in line.split(",")]
            if len(parts) < 7:
                continue
            try:
                sm.append(float(parts[0]))
                mx.append(float(parts[1]))
            except ValueError:
                continue
            for nm, v in zip(names, parts[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(nm)
        self.f.close()
        os.unlink(self.f.name)
        if sm:
            out.update(sm_mhz=statistics.median(sm), sm_max_mhz=max(mx), reasons=sorted(reasons), samples=len(sm))
        return out


# ---- synthetic JRDB-shaped input -----------------------------------------------------------------
def make_frames(num, seed0=1994):
    """`num` distinct seeded frames of configs[1] -> xyz (num,N,3), feat (num,N,1), boxes (num,M,7) float32."""
    import numpy as np
    from multimodal_gar_b200 import synthetic
    xs, fs, bs = [], [], []
    for i in range(num):
        pts, boxes = synthetic.make_frame(seed0 + i, N=CFG["N"], M=CFG["M"])
        xs.append(pts[:, :3])
        fs.append(pts[:, 3:4])
        bs.append(boxes)
    return (np.ascontiguousarray(np.stack(xs)), np.ascontiguousarray(np.stack(fs)), np.ascontiguousarray(np.stack(bs)))


def make_encoder():
    """Random-init weights of the configs[1] architecture (seeded: both arms use the same numbers)."""
    import torch
    from multimodal_gar_b200 import lidar_branch
    torch.manual_seed(1994)
    enc = lidar_branch.PointPersonEncoder(CFG["C_in"], CFG["S"], CFG["P"], CFG["radius"], CFG["K"], CFG["mlp"])
    return enc.eval()


# ---- the path on the host cores (oracle port; checker code, timed only as the reported baseline) ----
_CPU_STATE = {}


def _cpu_init(threads):
    import torch
    torch.set_num_threads(threads)
    from oracle import oracle as O
    O.lib()
    _CPU_STATE["mlp"] = make_encoder().sa.mlps[0]
    _CPU_STATE["O"] = O


def cpu_frame(args):
    """One frame through the oracle: points_in_boxes (first hit), roipoint pooling, canonical frame,
    FPS, ball query, grouping, shared MLP (torch CPU, the reference's own nn.Sequential), max."""
    import numpy as np
    xyz, feat, boxes = args
    O, mlp = _CPU_STATE["O"], _CPU_STATE["mlp"]
    S, P, K, r, M = CFG["S"], CFG["P"], CFG["K"], CFG["radius"], boxes.shape[0]
    t0 = time.perf_counter()
    O.points_in_boxes_gpu(xyz[None], boxes[None])
    t1 = time.perf_counter()
    pooled, empty = O.roipoint_pool3d(xyz[None], feat[None], boxes[None], (0.0, 0.0, 0.0), S)
    pooled = O.canonical_transform(pooled, boxes[None], empty)
    t2 = time.perf_counter()
    cxyz = np.ascontiguousarray(pooled[..., :3])
    fps = O.farthest_point_sample(cxyz, P)
    t3 = time.perf_counter()
    new_xyz = np.take_along_axis(cxyz, fps[..., None].astype(np.int64), axis=1).reshape(-1, 3)
    cnt, ncnt = np.full(M, S, np.int32), np.full(M, P, np.int32)
    cf = np.ascontiguousarray(pooled[..., 3:]).reshape(M * S, -1)
    nf, _ = O.query_and_group(r, K, cxyz.reshape(-1, 3), cnt, new_xyz, ncnt, cf)
    t4 = time.perf_counter()
    outs = [O.sa_mlp_forward(mlp, nf[i:i + 4096]) for i in range(0, nf.shape[0], 4096)]
    person = np.concatenate(outs).reshape(M, P, -1).max(axis=1)
    person[empty.reshape(-1) > 0] = 0
    t5 = time.perf_counter()
    return float(person.sum()), (t1 - t0, t2 - t1, t3 - t2, t4 - t3, t5 - t4)


def usable_cores():
    try:
        return len(os.sched_getaffinity(0))
    except AttributeError:
        return os.cpu_count() or 1


def cpu_baseline_inline(frames, n_frames):
    """Scalar port on ONE host thread over a bounded sample of the workload (rank 0, N=1)."""
    _cpu_init(1)
    xyz, feat, boxes = frames
    t0 = time.perf_counter()
    stage = [0.0] * 5
    for i in range(n_frames):
        _, st = cpu_frame((xyz[i], feat[i], boxes[i]))
        stage = [a + b for a, b in zip(stage, st)]
    dt = time.perf_counter() - t0
    names = ["points_in_boxes", "roipoint_pool", "fps", "ball_query+group", "sa_mlp(torch-cpu)"]
    return dict(value=n_frames / dt, unit="frames/s", cores=1, kind="port",
                sample="%d configs[1] frame(s), oracle C port + torch-CPU shared MLP, 1 thread, %.1f s" % (n_frames, dt),
                stage_seconds_per_frame={k: round(v / n_frames, 4) for k, v in zip(names, stage)})


# ---- reference arm: the path on the box's host cores, all the threads it can use -----------------
def run_reference(args, rank, world):
    if rank != 0:
        return
    from concurrent.futures import ProcessPoolExecutor
    import multiprocessing as mp
    cores = usable_cores()
    workers = max(1, min(cores, 32))                  # one frame per worker, 1 thread each; bounded memory
    frames = make_frames(workers)
    jobs = [(frames[0][i], frames[1][i], frames[2][i]) for i in range(workers)]
    with ProcessPoolExecutor(workers, mp_context=mp.get_context("spawn"), initializer=_cpu_init, initargs=(1,)) as ex:
        for _ in range(max(1, args.warmup)):          # also starts the workers (imports torch in each)
            list(ex.map(cpu_frame, jobs))
        t0 = time.perf_counter()
        for _ in range(args.steps):
            list(ex.map(cpu_frame, jobs))
        dt = time.perf_counter() - t0
    value = workers * args.steps / dt
    sample = "%d frames per step (one per worker process, 1 thread each), oracle C port + torch-CPU shared MLP" % workers
    line = dict(metric=METRIC, value=value, unit="frames/s", n_gpus=args.gpus, steps=args.steps, warmup=args.warmup,
                ms_per_step=1e3 * dt / args.steps, higher_is_better=True, scaling="weak", vs_baseline=None,
                dtype="f32", data="synthetic", impl="reference",
                config=dict(workload="configs[1]: 65536 pts x 40 boxes, S=512 P=512 r=0.3 K=32 MLP 4-64-128",
                            frames_per_step=workers),
                cpu_baseline=dict(value=value, unit="frames/s", cores=workers, kind="port", sample=sample,
                                  host_cores=cores),
                e2e=dict(value=value, unit="frames/s", h2d_bytes_per_step=0, d2h_bytes_per_step=0))
    print(json.dumps(line), flush=True)


# ---- native arm ---------------------------------------------------------------------------------------
def stage_work(F, keep_sa=True):
    """Algorithmic bytes (SURVEY.md 8d per-stage figures) and flops of one step of F frames, per stage."""
    N, M, S, P, K, C2 = CFG["N"], CFG["M"], CFG["S"], CFG["P"], CFG["K"], CFG["mlp"][-1]
    sa_out = 4 * M * C2 * (P if keep_sa else 1)
    return {
        "points_in_boxes": dict(bytes=F * (12 * N + 28 * M + 4 * N), flops=0),
        "roipoint_pool": dict(bytes=F * (16 * N + 28 * M + 16 * M * S + 4 * M), flops=0),
        "fps": dict(bytes=F * (12 * M * S + 4 * M * P), flops=0),
        "ball_query": dict(bytes=F * (12 * M * S + 12 * M * P + 4 * M * P * K), flops=0),
        "fps_ball_query": dict(bytes=F * (12 * M * S + 4 * M * P + 12 * M * P + 4 * M * P * K), flops=0),   # fused: one read of the clouds
        "sa_mlp": dict(bytes=F * (16 * M * S + 12 * M * P + 4 * M * P * K + sa_out), flops=F * mlp_flops_per_frame()),
        "person_reduce": dict(bytes=F * (4 * M * P * C2 + 4 * M * C2), flops=0),
    }


def build_batches(F, rank, dev):
    """Two alternating device batches (inputs 2 x F MB > the 126 MB L2 for F = 64; each step also writes
    ~14 MB/frame of stage outputs) built from 16 distinct seeded frames, points re-permuted per slot."""
    import numpy as np
    import torch
    base = make_frames(min(F, 16), seed0=1994 + 100 * rank)
    g = torch.Generator().manual_seed(2023 + rank)
    host, devb = [], []
    for which in range(2):
        xs, fs, bs = [], [], []
        for i in range(F):
            j = i % base[0].shape[0]
            perm = torch.randperm(CFG["N"], generator=g).numpy() if (i >= base[0].shape[0] or which) else np.arange(CFG["N"])
            xs.append(base[0][j][perm])
            fs.append(base[1][j][perm])
            bs.append(base[2][j])
        h = [torch.from_numpy(np.ascontiguousarray(np.stack(a))).pin_memory() for a in (xs, fs, bs)]
        host.append(h)
        devb.append([t.to(dev) for t in h])
    return host, devb


def run_native(args, rank, world, local_rank):
    import torch
    import torch.distributed as dist
    from multimodal_gar_b200 import _lib
    if not torch.cuda.is_available():
        raise SystemExit("bench.py (native arm) needs a CUDA device: there is no CPU fallback")
    _lib.lib()
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    if world > 1:
        os.environ.setdefault("NCCL_DEBUG_FILE", "/dev/stderr")     # NCCL's version banner must not share stdout with the JSON line
        dist.init_process_group("nccl", device_id=dev)
    F = args.frames_per_step
    enc = make_encoder().to(dev)
    host, devb = build_batches(F, rank, dev)
    work = stage_work(F, enc.keep_sa_features)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def max_over_ranks(ms):
        if world > 1:
            t = torch.tensor([ms], dtype=torch.float64, device=dev)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            return float(t.item())
        return ms

    # ---- device-resident: `value` -------------------------------------------------------------------
    with torch.no_grad():
        for i in range(args.warmup):
            enc(*devb[i & 1])
        barrier()
        sampler = ClockSampler(local_rank) if rank == 0 else None
        marks = []
        n0 = _lib.launch_count()
        ev0 = torch.cuda.Event(enable_timing=True)
        ev0.record()
        for i in range(args.steps):
            evs = [("start", torch.cuda.Event(enable_timing=True))]
            evs[0][1].record()

            def mark(name, evs=evs):
                e = torch.cuda.Event(enable_timing=True)
                e.record()
                evs.append((name, e))
            enc(*devb[i & 1], mark=mark)
            marks.append(evs)
        ev1 = torch.cuda.Event(enable_timing=True)
        ev1.record()
        barrier()
        launches = _lib.launch_count() - n0
        ms_total = max_over_ranks(ev0.elapsed_time(ev1))
        clocks = sampler.stop() if sampler else None
    stage_ms = {}
    for evs in marks:
        for (_, a), (name, b) in zip(evs, evs[1:]):
            stage_ms[name] = stage_ms.get(name, 0.0) + a.elapsed_time(b) / len(marks)

    # ---- end to end through the public API with HOST buffers: `e2e` ---------------------------------
    # The public host-buffer API (multimodal_gar_b200.pipeline.HostFramePipeline): every step copies its pinned host
    # inputs to the device, runs the branch and copies the person features back; the three legs run on their own
    # streams, so the copy of step i+1 overlaps the kernels of step i.  All of it is inside the timed region.
    from multimodal_gar_b200 import pipeline
    out_host = [torch.empty((F, CFG["M"], CFG["mlp"][-1]), dtype=torch.float32).pin_memory() for _ in range(2)]
    pipe = pipeline.HostFramePipeline(enc, dev, depth=2)
    with torch.no_grad():
        for i in range(max(2, args.warmup // 2)):
            pipe.submit(*host[i & 1], out_host[i & 1])
        pipe.synchronize()
        barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for i in range(args.steps):
            pipe.submit(*host[i & 1], out_host[i & 1])
        pipe.join()
        e1.record()
        pipe.synchronize()
        barrier()
        ms_e2e = max_over_ranks(e0.elapsed_time(e1))
        # single-frame latency (F = 1), reported beside the throughput number
        one = [t[:1].contiguous() for t in devb[0]]
        for _ in range(5):
            enc(*one)
        torch.cuda.synchronize()
        l0, l1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        l0.record()
        for _ in range(20):
            enc(*one)
        l1.record()
        torch.cuda.synchronize()
        lat_ms = l0.elapsed_time(l1) / 20
    h2d = sum(t.numel() * t.element_size() for t in host[0])
    d2h = out_host[0].numel() * out_host[0].element_size()
    return dict(keep_sa=enc.keep_sa_features, F=F, ms_total=ms_total, ms_e2e=ms_e2e, stage_ms=stage_ms, launches=launches, clocks=clocks,
                h2d=h2d, d2h=d2h, lat_ms=lat_ms, work=work, host=host)


def main():
    args = parse_args()
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
    if args.impl == "reference":
        run_reference(args, rank, world)
        return
    r = run_native(args, rank, world, local_rank)
    import torch.distributed as dist
    if rank == 0:
        pk = peaks()
        F, K = r["F"], args.steps
        frames_total = F * K * world
        value = frames_total / (r["ms_total"] * 1e-3)
        stages = {}
        for name, ms in r["stage_ms"].items():
            w = r["work"].get(name, dict(bytes=0, flops=0))
            stages[name] = dict(ms=round(ms, 4), algorithmic_GBps=round(w["bytes"] / (ms * 1e-3) / 1e9, 2),
                                TFLOPs=round(w["flops"] / (ms * 1e-3) / 1e12, 3))
        dom = max(r["stage_ms"], key=r["stage_ms"].get)
        w, ms = r["work"][dom], r["stage_ms"][dom]
        if w["flops"] > 0:        # the shared MLP: the one dense contraction -> tensor-pipe roofline
            ach = w["flops"] / (ms * 1e-3) / 1e12
            roof = dict(bound="tensor", achieved=ach, peak=pk["bf16_sustained"], unit="TFLOP/s",
                        frac=ach / pk["bf16_sustained"], traffic=ncu_traffic("sa_mlp", F), kernel=dom,
                        note="algorithmic fp32 FLOPs (2*M*P*K*8448 per frame) over the kernel's CUDA-event time, against the "
                             "%s sustained bf16 tensor peak; the kernel issues 3 fp16 MMAs per product (3xFP16 split, "
                             "fp32-faithful), i.e. the tensor pipe does 3x the algorithmic work (ncu: pipe 82 %% busy, "
                             "profiles/r01_full_summary_final.md) - frac compares fp32-faithful output with a bf16 library "
                             "GEMM; traffic = ncu dram bytes per launch (profiles/r01_traffic.json)" % pk["source"])
        else:
            ach = w["bytes"] / (ms * 1e-3) / 1e9
            roof = dict(bound="hbm", achieved=ach, peak=pk["hbm"], unit="GB/s", frac=ach / pk["hbm"],
                        traffic=ncu_traffic(dom, F), kernel=dom, note="of %s" % pk["source"])
        roof["share_of_step"] = ms / (r["ms_total"] / K)
        pipeline_GBps = algorithmic_bytes_per_frame(keep_sa=r["keep_sa"]) * value / 1e9
        line = dict(metric=METRIC, value=value, unit="frames/s", n_gpus=world, steps=K, warmup=args.warmup,
                    ms_per_step=r["ms_total"] / K, higher_is_better=True, scaling="weak", vs_baseline=None,
                    dtype="f32", data="synthetic",
                    config=dict(workload="configs[1]: 65536 pts x 40 boxes, S=512 P=512 r=0.3 K=32 MLP 4-64-128",
                                frames_per_step=F * world, frames_per_step_per_gpu=F,
                                l2="two alternating input batches (2 x %d MB) + ~%d MB of stage outputs per step > 126 MB L2"
                                   % (F, F * (14 if r["keep_sa"] else 4)),
                                sa_features="stored" if r["keep_sa"] else "kept on chip (person max fused into the MLP kernel)",
                                weights="random-init (seed 1994), eval-mode BatchNorm folded"),
                    e2e=dict(value=frames_total / (r["ms_e2e"] * 1e-3), unit="frames/s",
                             h2d_bytes_per_step=r["h2d"], d2h_bytes_per_step=r["d2h"]),
                    gpu_launches=r["launches"], clocks=r["clocks"], roofline=roof, stages=stages,
                    pipeline_algorithmic_GBps=pipeline_GBps, pipeline_hbm_frac=pipeline_GBps / pk["hbm"] / world,
                    latency_1frame_ms=r["lat_ms"])
        if world == 1 and not args.no_cpu_baseline:
            n = args.cpu_frames or 8
            import numpy as np
            frames = tuple(np.ascontiguousarray(t[:n].numpy()) for t in r["host"][0])
            line["cpu_baseline"] = cpu_baseline_inline(frames, min(n, F))
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
