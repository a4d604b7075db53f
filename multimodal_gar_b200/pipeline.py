"""Host-buffer front end of the LiDAR branch: frames arrive in (pinned) host memory, person features leave to
host memory, and the three legs - H2D copy, the five kernels, D2H copy - run on their own CUDA streams so the copy
of batch i+1 overlaps the compute of batch i.

This is what a dataloader-fed caller uses (the reference feeds `LiDAR_Backbone` from a torch DataLoader and
`.cuda()`s every batch on the compute stream, train_func.py / gat_model.py:910): same results as calling the
encoder on device tensors, frame order preserved, no extra copies (inputs land in per-slot device buffers that the
kernels read in place).
"""
import ctypes
import weakref

import torch

from . import _lib


def pinned_empty(shape, dtype=torch.float32, write_combined=False):
    """A page-locked host tensor from the library's allocator (cudaHostAlloc, portable): what a dataloader writes its frames
    into.  write_combined=True asks for write-combined memory - the CPU must only WRITE it (reads are uncached and slow);
    the copy engines read it without snooping the CPU caches, which can raise host-to-device throughput when several GPUs
    pull from one host.  Freed when the tensor is garbage-collected."""
    n = 1
    for d in shape:
        n *= int(d)
    nbytes = max(1, n * torch.empty((), dtype=dtype).element_size())
    p = ctypes.c_void_p()
    _lib.check(_lib.lib().mgar_host_alloc(ctypes.addressof(p), nbytes, 1 if write_combined else 0), "host_alloc")
    buf = (ctypes.c_char * nbytes).from_address(p.value)
    t = torch.frombuffer(buf, dtype=dtype, count=n).view(*shape)
    weakref.finalize(buf, lambda addr=p.value: _lib.lib().mgar_host_free(addr))
    return t


class HostFramePipeline:
    """`submit(xyz_h, feat_h, boxes_h, out_h)` enqueues one batch; `synchronize()` waits for everything.

    xyz_h (B,N,3), feat_h (B,N,C), boxes_h (B,M,7) float32 host tensors (pinned for asynchronous copies),
    out_h (B,M,C_out) float32 host tensor that receives the person features.  `depth` batches may be in flight;
    slot buffers are allocated on first use and reused while the shapes stay the same."""

    def __init__(self, encoder, device, depth=2):
        self.enc = encoder
        self.dev = torch.device(device)
        self.depth = depth
        self.s_in = torch.cuda.Stream(self.dev)
        self.s_run = torch.cuda.Stream(self.dev)
        self.s_out = torch.cuda.Stream(self.dev)
        self.slots = [None] * depth
        self.n = 0

    def _slot(self, k, tensors):
        s = self.slots[k]
        if s is None or any(a.shape != b.shape for a, b in zip(s["dev"], tensors)):
            if s is not None:
                # the kernels / copies still using the old buffers must be done before the allocator may reuse them
                for ev in (s["consumed"], s["flushed"]):
                    if ev is not None:
                        ev.synchronize()
            # allocated ON the copy stream (the first writer), and announced to the stream whose kernels read them,
            # so the caching allocator never hands the block to pending work of another stream
            with torch.cuda.stream(self.s_in):
                dev = [torch.empty(t.shape, dtype=t.dtype, device=self.dev) for t in tensors]
            for d in dev:
                d.record_stream(self.s_run)
            s = dict(dev=dev, ready=torch.cuda.Event(), consumed=None, flushed=None, out=None)
            self.slots[k] = s
        return s

    @torch.no_grad()
    def submit(self, xyz_h, feat_h, boxes_h, out_h):
        k = self.n % self.depth
        self.n += 1
        s = self._slot(k, (xyz_h, feat_h, boxes_h))
        with torch.cuda.stream(self.s_in):
            if s["consumed"] is not None:
                self.s_in.wait_event(s["consumed"])          # the kernels of the batch that used this slot are done
            for d, h in zip(s["dev"], (xyz_h, feat_h, boxes_h)):
                d.copy_(h, non_blocking=True)
            s["ready"].record(self.s_in)
        with torch.cuda.stream(self.s_run):
            self.s_run.wait_event(s["ready"])
            if s["flushed"] is not None:
                self.s_run.wait_event(s["flushed"])          # the previous result of this slot has left the device
            person = self.enc(*s["dev"])
            s["consumed"] = torch.cuda.Event()
            s["consumed"].record(self.s_run)
        with torch.cuda.stream(self.s_out):
            self.s_out.wait_event(s["consumed"])
            person.record_stream(self.s_out)
            out_h.copy_(person, non_blocking=True)
            s["flushed"] = torch.cuda.Event()
            s["flushed"].record(self.s_out)
        s["out"] = person
        return k

    def join(self, stream=None):
        """Make `stream` (default: the current stream) wait for every enqueued leg - for timing with events."""
        stream = stream or torch.cuda.current_stream(self.dev)
        for st in (self.s_in, self.s_run, self.s_out):
            ev = torch.cuda.Event()
            ev.record(st)
            stream.wait_event(ev)

    def synchronize(self):
        for st in (self.s_in, self.s_run, self.s_out):
            st.synchronize()
