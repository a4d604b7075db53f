"""CUDA-graph replay of the LiDAR branch: the whole step (8 launches of libmgar_lidar.so + the memsets and small tensor
ops around them, the two side-by-side sampling kernels with their fork / join included) captured once and replayed with one
`cudaGraphLaunch` - for serving loops whose step is short enough for launch gaps to show (a single frame: 0.2 ms).

The library allocates nothing and synchronises nothing (include/mgar_lidar.h), so its calls are capturable as they are; the
scratch tensors the Python layer takes from torch's caching allocator come from the graph's private pool during capture.
Inputs are copied into static buffers before a replay, the result is a static buffer (clone it to keep it).  A graph holds
the device addresses of the encoder's folded parameters: capture again after changing weights or running statistics."""
import torch


class GraphedPersonEncoder:
    """`PointPersonEncoder.forward(points, point_features, boxes)` for ONE fixed shape, replayed from a CUDA graph.

        g = GraphedPersonEncoder(enc, points, point_features, boxes)      # warm-up + capture (eval mode, no autograd)
        person = g(points2, point_features2, boxes2)                      # (B, M, C_out), a view of the static output

    The result of a replay is overwritten by the next one."""

    def __init__(self, encoder, points, point_features, boxes, warmup=3):
        if not (points.is_cuda and point_features.is_cuda and boxes.is_cuda):
            raise ValueError("GraphedPersonEncoder needs CUDA tensors")
        if encoder.training:
            raise ValueError("GraphedPersonEncoder captures the eval path: call encoder.eval() first")
        self.encoder = encoder
        self.static_in = [points.clone(), point_features.clone(), boxes.clone()]
        self.graph = torch.cuda.CUDAGraph()
        side = torch.cuda.Stream(device=points.device)
        side.wait_stream(torch.cuda.current_stream(points.device))
        with torch.no_grad(), torch.cuda.stream(side):
            for _ in range(max(1, warmup)):                 # lazy state (folded parameters, the library's side stream) settles here
                self.encoder(*self.static_in)
        torch.cuda.current_stream(points.device).wait_stream(side)
        torch.cuda.synchronize(points.device)
        with torch.no_grad(), torch.cuda.graph(self.graph):
            self.static_out = self.encoder(*self.static_in)

    def __call__(self, points, point_features, boxes):
        for dst, src in zip(self.static_in, (points, point_features, boxes)):
            if dst.shape != src.shape or dst.dtype != src.dtype:
                raise ValueError("GraphedPersonEncoder was captured for shapes %s" % [tuple(t.shape) for t in self.static_in])
            if dst.data_ptr() != src.data_ptr():
                dst.copy_(src, non_blocking=True)
        self.graph.replay()
        return self.static_out


class GraphedCallable:
    """Any zero-argument eval-mode computation over tensors that stay where they are (`fn` closes over them), captured once:
    `g = GraphedCallable(fn, device); out = g()` replays it.  Used by bench.py for configs[2] (LiDAR branch + DAFM inputs +
    fusion): with 8 frames per rank the ~60 launches of that step are what it lasts."""

    def __init__(self, fn, device, warmup=3):
        self.graph = torch.cuda.CUDAGraph()
        side = torch.cuda.Stream(device=device)
        side.wait_stream(torch.cuda.current_stream(device))
        with torch.no_grad(), torch.cuda.stream(side):
            for _ in range(max(1, warmup)):
                fn()
        torch.cuda.current_stream(device).wait_stream(side)
        torch.cuda.synchronize(device)
        with torch.no_grad(), torch.cuda.graph(self.graph):
            self.out = fn()

    def __call__(self):
        self.graph.replay()
        return self.out
