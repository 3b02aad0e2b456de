"""Host-facing streaming front end: batches of frame pairs go host -> device -> host with the copies of neighbouring
batches hidden behind the matcher (the reference's video loops hand frames to cv2 one pair at a time from host memory,
stereo_calibration/scripts/ball_drop.py:370-420; this is the same hand-off for a GPU that wants many pairs in flight).

Three CUDA streams: copy-in, compute (the caller's current stream), copy-out; two device slots per direction.  While
batch i is being matched, batch i+1 is already on its way in and the disparities / box medians of batch i-1 are on
their way out.  All ordering is by CUDA events; nothing here synchronises the host except `synchronize()`.
"""
from __future__ import annotations

import numpy as np
import torch

from . import api


class HostPipeline:
    def __init__(self, matcher: "api.StereoSGBM", device=None, slots: int = 2):
        self.m = matcher
        self.dev = torch.cuda.current_device() if device is None else (torch.device(device).index or 0)
        self.slots = int(slots)
        with torch.cuda.device(self.dev):
            self.s_in = torch.cuda.Stream()
            self.s_out = torch.cuda.Stream()
            self.ev_in = [torch.cuda.Event() for _ in range(self.slots)]
            self.ev_comp = [torch.cuda.Event() for _ in range(self.slots)]
            self.ev_out = [torch.cuda.Event() for _ in range(self.slots)]
        self.L = [None] * self.slots
        self.R = [None] * self.slots
        self.D = [None] * self.slots
        self.i = 0
        self.used = [False] * self.slots

    @staticmethod
    def _host(t, what):
        if isinstance(t, np.ndarray):
            t = torch.from_numpy(np.ascontiguousarray(t))
        if not isinstance(t, torch.Tensor) or t.is_cuda:
            raise ValueError(f"{what}: host tensor / array expected")
        return t

    def submit(self, left_h, right_h, disp_h, Q=None, boxes=None, box_image=None, med_h=None, minDisparity=None):
        """Enqueue one batch: left_h/right_h uint8 [B,H,W(,3)] host (pinned for true overlap), disp_h int16 [B,H,W] host
        output; with Q + boxes (device or host, N x 4) also the per-box median depth into med_h (float32 [N] host).
        Returns immediately; the outputs are valid after `synchronize()` (or after an event recorded behind `flush()`)."""
        left_h, right_h, disp_h = self._host(left_h, "left"), self._host(right_h, "right"), self._host(disp_h, "disp")
        if left_h.shape != right_h.shape or left_h.dtype != torch.uint8 or disp_h.dtype != torch.int16:
            raise ValueError("left/right must be equal-shaped uint8 batches, disp int16")
        s = self.i % self.slots
        self.i += 1
        d = f"cuda:{self.dev}"
        with torch.cuda.device(self.dev):
            main = torch.cuda.current_stream()
            if self.L[s] is None or self.L[s].shape != left_h.shape:
                if self.used[s]:
                    # the old slot buffers may still be read by the copy-out stream / written by the copy-in stream:
                    # let the allocating stream wait for both before the caching allocator can reuse that memory
                    main.wait_event(self.ev_out[s])
                    main.wait_event(self.ev_in[s])
                    self.used[s] = False
                self.L[s] = torch.empty(left_h.shape, dtype=torch.uint8, device=d)
                self.R[s] = torch.empty(left_h.shape, dtype=torch.uint8, device=d)
                self.D[s] = torch.empty(disp_h.shape, dtype=torch.int16, device=d)
            if self.used[s]:
                self.s_in.wait_event(self.ev_comp[s])        # the matcher is done reading this input slot
            with torch.cuda.stream(self.s_in):
                self.L[s].copy_(left_h, non_blocking=True)
                self.R[s].copy_(right_h, non_blocking=True)
                self.ev_in[s].record(self.s_in)
            main.wait_event(self.ev_in[s])
            if self.used[s]:
                main.wait_event(self.ev_out[s])              # the previous result of this slot has left the device
            self.m._run(self.L[s], self.R[s], self.dev, out=self.D[s])
            med = None
            if boxes is not None:
                if Q is None:
                    raise ValueError("box medians need Q")
                md = self.m.getMinDisparity() if minDisparity is None else minDisparity
                med = api.box_median_depth(self.D[s], Q, boxes, minDisparity=md, box_image=box_image)
            self.ev_comp[s].record(main)
            self.s_out.wait_event(self.ev_comp[s])
            with torch.cuda.stream(self.s_out):
                disp_h.copy_(self.D[s], non_blocking=True)
                if med is not None and med_h is not None:
                    self._host(med_h, "med").copy_(med, non_blocking=True)
                    med.record_stream(self.s_out)
                self.ev_out[s].record(self.s_out)
            self.used[s] = True

    def flush(self):
        """Make the caller's current stream wait for every copy enqueued so far (an event recorded on it afterwards
        covers the whole pipeline; no host synchronisation)."""
        with torch.cuda.device(self.dev):
            main = torch.cuda.current_stream()
            for s in range(self.slots):
                if self.used[s]:
                    main.wait_event(self.ev_out[s])

    def synchronize(self):
        with torch.cuda.device(self.dev):
            self.s_out.synchronize()
            torch.cuda.current_stream().synchronize()
