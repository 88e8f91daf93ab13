"""Host <-> device staging around an engine for a video stream.

The engines take device tensors (like the reference's, whose evaluator calls ``.cuda()`` on every frame,
networks/managers/evaluator.py:385-390).  A host loop that copies frame f, runs it, and reads its mask back
before touching frame f+1 leaves the GPU idle during both copies.  ``FrameStreamer`` double-buffers both
directions with two copy streams: the upload of frame f+1 starts as soon as the host asks for it (three device
buffers, so it normally lands while frame f-1 is still computing), and the label map of frame f is read back
while frame f+1 computes (results are consumed one frame late).  Every frame is still
uploaded from pinned host memory and every result is still read on the host.
"""
import collections

import torch


class FrameStreamer:
    def __init__(self, device, depth=3):
        self.device = torch.device(device)
        self.depth = depth
        self._up = torch.cuda.Stream(device=self.device)
        self._down = torch.cuda.Stream(device=self.device)
        self._bufs = [None] * depth           # device image buffers
        self._free = [None] * depth           # event: the compute stream is done with buffer i
        self._staged = collections.deque()    # (buffer index, upload-done event)
        self._next = 0
        self._results = collections.deque()   # (pinned host tensor, device source kept alive, done event)
        self._host = [None] * (depth + 1)
        self._host_next = 0

    def preallocate(self, image_shape, result_shape, image_dtype=torch.float32, result_dtype=torch.uint8):
        """Allocate the device image buffers and the pinned result buffers now (pinning host memory costs
        milliseconds; do it before the stream starts, not on the first frames)."""
        for i in range(self.depth):
            self._bufs[i] = torch.empty(image_shape, dtype=image_dtype, device=self.device)
        for j in range(len(self._host)):
            self._host[j] = torch.empty(result_shape, dtype=result_dtype).pin_memory()
        return self

    # ---------------------------------------------------------------- uploads
    def stage(self, host_img):
        """Start the upload of a pinned host image; returns immediately."""
        i = self._next
        self._next = (i + 1) % self.depth
        if self._bufs[i] is None or self._bufs[i].shape != host_img.shape:
            self._bufs[i] = torch.empty(host_img.shape, dtype=host_img.dtype, device=self.device)
        if self._free[i] is not None:
            self._up.wait_event(self._free[i])            # the frame that last used this buffer has been consumed
        with torch.cuda.stream(self._up):
            self._bufs[i].copy_(host_img, non_blocking=True)
            ev = torch.cuda.Event()
            ev.record()
        self._staged.append((i, ev))

    def stage_stack(self, host_imgs):
        """Start the upload of several pinned host images into ONE device buffer [n, ...] (the stack an engine takes
        as ``next_img``: the frames of its next n calls); returns immediately."""
        i = self._next
        self._next = (i + 1) % self.depth
        shape = (len(host_imgs) * host_imgs[0].shape[0],) + tuple(host_imgs[0].shape[1:])
        if self._bufs[i] is None or tuple(self._bufs[i].shape) != shape or self._bufs[i].dtype != host_imgs[0].dtype:
            self._bufs[i] = torch.empty(shape, dtype=host_imgs[0].dtype, device=self.device)
        if self._free[i] is not None:
            self._up.wait_event(self._free[i])
        n0 = host_imgs[0].shape[0]
        with torch.cuda.stream(self._up):
            for j, im in enumerate(host_imgs):
                self._bufs[i][j * n0:(j + 1) * n0].copy_(im, non_blocking=True)
            ev = torch.cuda.Event()
            ev.record()
        self._staged.append((i, ev))

    def acquire(self):
        """Oldest staged image as a device tensor; the current stream waits for its upload."""
        i, ev = self._staged.popleft()
        torch.cuda.current_stream().wait_event(ev)
        self._cur = i
        return self._bufs[i]

    def release(self, tensor=None):
        """Call after the work reading an acquired image has been enqueued on (or is awaited by) the current
        stream; ``tensor`` names the image (default: the one acquired last)."""
        idx = self._cur
        if tensor is not None:
            idx = next(i for i, b in enumerate(self._bufs)
                       if b is tensor or (b is not None and tensor.data_ptr() == b.data_ptr()))
        ev = torch.cuda.Event()
        ev.record()
        self._free[idx] = ev

    # ---------------------------------------------------------------- results
    def push_result(self, dev_tensor):
        """Start the read-back of a device tensor (e.g. the uint8 label map) into pinned host memory."""
        j = self._host_next
        self._host_next = (j + 1) % len(self._host)
        if self._host[j] is None or self._host[j].shape != dev_tensor.shape or self._host[j].dtype != dev_tensor.dtype:
            self._host[j] = torch.empty(dev_tensor.shape, dtype=dev_tensor.dtype).pin_memory()
        ready = torch.cuda.Event()
        ready.record()
        self._down.wait_event(ready)
        dev_tensor.record_stream(self._down)
        with torch.cuda.stream(self._down):
            self._host[j].copy_(dev_tensor, non_blocking=True)
            done = torch.cuda.Event()
            done.record()
        self._results.append((self._host[j], dev_tensor, done))

    def pending_results(self):
        return len(self._results)

    def pop_result(self):
        """Oldest result as a pinned host tensor (blocks until its copy has landed)."""
        host, _, done = self._results.popleft()
        done.synchronize()
        return host
