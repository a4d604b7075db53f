"""Round latency of the fused FPS + ball query kernel: b clouds of 512 distinct points (512 rounds each), one CTA per SM or
fewer -> kernel time / 511 rounds = the dependent chain of one round, per geometry (MGAR_FPS_BQ_GEOM / MGAR_FPS_BQ)."""
import json
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
from multimodal_gar_b200 import fastpath  # noqa: E402

dev = torch.device("cuda:0")
res = {}
for b in (1, 148, 148 * 4, 148 * 10):
    rng = np.random.default_rng(b)
    c4 = torch.from_numpy(rng.uniform(-1, 1, (b, 512, 4)).astype(np.float32)).to(dev)
    x = c4[:, :, 0:3]
    for geom in ("2,8", "4,4", "8,2", "1,16", "warp"):
        os.environ.pop("MGAR_FPS_BQ", None)
        os.environ.pop("MGAR_FPS_BQ_GEOM", None)
        if geom == "warp":
            if b < 592:
                continue
            os.environ["MGAR_FPS_BQ"] = "warp"
        else:
            os.environ["MGAR_FPS_BQ_GEOM"] = geom
        try:
            for _ in range(2):
                fastpath.fps_ball_query(x, 512, 0.3, 32)
        except Exception as e:
            res["%d %s" % (b, geom)] = "n/a"
            continue
        torch.cuda.synchronize()
        a, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        for _ in range(5):
            fastpath.fps_ball_query(x, 512, 0.3, 32)
        e.record()
        torch.cuda.synchronize()
        ms = a.elapsed_time(e) / 5
        res["%d %s" % (b, geom)] = dict(ms=round(ms, 4), cycles_per_round=round(ms * 1e-3 * 1.965e9 / 511))
print(json.dumps(res, indent=1))
