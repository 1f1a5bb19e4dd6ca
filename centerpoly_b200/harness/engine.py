"""Inference engine around the hot path: DLA-34 polydet network (harness/dla34.py) -> ``hm.sigmoid_()`` ->
``polydet_decode`` (+ the coordinate back-mapping of ``polydet_post_process``, fused into the decode), i.e. what
``PolydetDetector.process`` and ``post_process`` do (src/lib/detectors/polydet.py:25-60), captured
once into a CUDA graph so the ~250 launches of a frame replay without host overhead.

``infer_resident`` replays on a frame already in HBM; ``infer_host`` is the end-to-end call: pinned host
image -> H2D -> graph -> D2H of the (B,K,2N+7) detections, double-buffered so the copy of frame i+1 overlaps
the compute of frame i (the reference's ``images.to(device)`` / ``dets.cpu()``, base_detector.py:134,
detectors/polydet.py:46).
"""
from __future__ import annotations

import contextlib

import torch

from .. import decode as cdecode
from .. import post_process as cpost
from .. import raster as craster
from .dla34 import PolydetDLA34, calibrate_random_init

HEADS = {"hm": 8, "poly": 64, "pseudo_depth": 1, "reg": 2}


class PolydetEngine:
    def __init__(self, height=1024, width=2048, batch=1, dtype="bf16", K=100, rep="cartesian", heads=None,
                 device="cuda:0", seed=317, calibrate=True, use_graph=True, n_slots=2, optimize=True, post_process=True,
                 rasterise=False, thresh=0.05, net=None):
        self.device = torch.device(device)
        self.dtype = dtype
        self.K, self.rep = K, rep
        self.heads = dict(heads or HEADS)
        self.shape = (batch, 3, height, width)
        torch.manual_seed(seed)
        torch.backends.cudnn.benchmark = True     # as the reference does (src/main.py:32): let cuDNN time its algorithms
        if net is None:
            net = PolydetDLA34(self.heads).to(self.device).to(memory_format=torch.channels_last)
            if calibrate:
                calibrate_random_init(net, height, width, seed=seed)
        else:       # a caller-built network (same module tree, e.g. loaded from a checkpoint); it is rewritten in place below
            net = net.to(self.device).to(memory_format=torch.channels_last)
        net = net.eval()
        self.fusion = None
        self.in_channels = 3
        if optimize:
            from .. import fuse
            # module rewrite; bf16 weights instead of per-call autocast casts
            self.fusion = fuse.optimize_for_inference(net, torch.bfloat16 if dtype == "bf16" else None)
            # (fuse.pad_stem_input_channels exists, but on this cuDNN the 8-channel bf16 stem is 2x SLOWER than the
            #  3-channel one - tools/stem_probe.py: 853 us vs 412 us - so the stem keeps its 3 input channels)
        self.optimize = optimize
        self._stem_direct = bool(self.fusion and self.fusion.get("stem"))
        self.net = net
        self.n_dcn_per_frame = sum(1 for m in net.modules() if hasattr(m, "conv_offset_mask"))
        if optimize:
            self._xin = torch.zeros((batch, self.in_channels, height, width), device=self.device,
                                    dtype=torch.bfloat16 if dtype == "bf16" else torch.float32
                                    ).contiguous(memory_format=torch.channels_last)
        # fused polydet_post_process (utils/post_process.py:105-112): frames are fed at network resolution, so the meta of
        # base_detector.py:55-57 (fix_res) is c = (W/2, H/2), s = max(H, W), out = (W/4, H/4); the (B,2,3) inverse affine
        # lives on the device and is read by the decode's gather kernel inside the captured graph
        self.post_affine = None
        if post_process:
            c = [[width / 2.0, height / 2.0]] * batch
            self.post_affine = torch.as_tensor(cpost.batch_inverse_affine(c, [float(max(height, width))] * batch, width // 4, height // 4),
                                               dtype=torch.float64).to(self.device)
        # optional: depth-ordered instance masks of format_and_write_to_cityscapes (datasets/dataset/cityscapes.py:226-279)
        # from the image-coordinate detections, inside the same captured graph (needs post_process)
        if rasterise and not post_process:
            raise ValueError("rasterise=True needs post_process=True (masks are drawn in image coordinates)")
        self.rasterise, self.thresh = bool(rasterise), float(thresh)
        self.stream = torch.cuda.Stream(self.device)
        self.copy_stream = torch.cuda.Stream(self.device)   # H2D
        self.d2h_stream = torch.cuda.Stream(self.device)    # D2H (separate, or it would serialise behind the next H2D)
        self.use_graph = use_graph
        n_poly = self.heads["poly"]
        self.slots = []
        for _ in range(n_slots):
            self.slots.append(dict(
                x=torch.zeros(self.shape, device=self.device).contiguous(memory_format=torch.channels_last),
                dets=torch.zeros((batch, K, n_poly + 7), device=self.device),
                host_out=torch.zeros((batch, K, n_poly + 7)).pin_memory(),
                masks=torch.zeros((batch, K, height, width), dtype=torch.uint8, device=self.device) if rasterise else None,
                raster=None, graph=None, h2d_done=torch.cuda.Event(), compute_done=torch.cuda.Event(), d2h_done=torch.cuda.Event()))
        self._warm_and_capture()

    # ------------------------------------------------------------------------------------------
    def _autocast(self):
        if self.dtype == "bf16" and not self.optimize:
            return torch.autocast("cuda", dtype=torch.bfloat16)
        return contextlib.nullcontext()

    def _heads(self, x):
        """The network part of a frame: head maps {hm (logits), poly, pseudo_depth, reg}."""
        if self.optimize and not self._stem_direct:
            # fp32 frame -> persistent network input in the model dtype: one conversion kernel
            self._xin[:, :3].copy_(x)
            x = self._xin
        # (with our stem kernel the fp32 channels_last frame is read directly and rounded to bf16 on load)
        return self.net(x)[-1]

    def heads_eager(self, slot_idx=0):
        """Head maps of the frame resident in a slot, computed eagerly on the engine stream (tests: the G3 gate)."""
        with torch.cuda.stream(self.stream), torch.no_grad(), self._autocast():
            out = {k: v.clone() for k, v in self._heads(self.slots[slot_idx]["x"]).items()}
        self.stream.synchronize()
        return out

    def _step(self, x):
        """One frame: network -> sigmoid -> decode (the body of PolydetDetector.process)."""
        with torch.no_grad(), self._autocast():
            out = self._heads(x)
            if self.optimize:   # the sigmoid of detectors/polydet.py:28 is fused into the decode kernel
                return cdecode.polydet_decode_from_logits(out["hm"], out["poly"], out["pseudo_depth"], reg=out.get("reg"),
                                                          K=self.K, rep=self.rep, post_affine=self.post_affine)
            hm = out["hm"].sigmoid_()
            return cdecode.polydet_decode(hm, out["poly"], out["pseudo_depth"], reg=out.get("reg"), K=self.K, rep=self.rep,
                                          post_affine=self.post_affine)

    def _frame(self, slot):
        """Everything one replay does for a slot: network + decode (+ back-mapping) into slot['dets'], then the masks."""
        slot["dets"].copy_(self._step(slot["x"]))
        if self.rasterise:
            _, _, h, w = self.shape
            slot["raster"] = craster.rasterise_instances(slot["dets"], w, h, self.thresh, out=slot["masks"])

    def _warm_and_capture(self):
        from .. import _lib
        with torch.cuda.stream(self.stream):
            for slot in self.slots:
                for _ in range(2):  # fills the packed-weight cache, cuDNN plans, allocator pools
                    before = sum(_lib.LAUNCHES.values())
                    self._frame(slot)
                    self.launches_per_step = sum(_lib.LAUNCHES.values()) - before   # our kernels in one warm frame
            self.stream.synchronize()
            if self.use_graph:
                for slot in self.slots:
                    g = torch.cuda.CUDAGraph()
                    with torch.cuda.graph(g, stream=self.stream):
                        self._frame(slot)
                    slot["graph"] = g
        self.stream.synchronize()

    def _run_slot(self, slot):
        if slot["graph"] is not None:
            slot["graph"].replay()
        else:
            self._frame(slot)

    # ------------------------------------------------------------------------------------------
    def set_resident_frames(self, frames):
        """frames: list of device tensors (B,3,H,W); copies them into the static slots."""
        for slot, f in zip(self.slots, frames):
            slot["x"].copy_(f)

    def infer_resident(self, slot_idx=0):
        """Enqueue one step on self.stream for the frame already resident in slot `slot_idx`."""
        with torch.cuda.stream(self.stream):
            self._run_slot(self.slots[slot_idx % len(self.slots)])

    def infer_host_stream(self, host_frames, n_steps):
        """End-to-end: n_steps frames taken round-robin from pinned `host_frames`; returns the last detections
        (host tensor).  H2D of step i+1 overlaps compute of step i; every step's detections are copied back."""
        nslots = len(self.slots)
        for i in range(n_steps):
            slot = self.slots[i % nslots]
            src = host_frames[i % len(host_frames)]
            with torch.cuda.stream(self.copy_stream):
                # the slot's previous compute must have consumed x before it is overwritten
                self.copy_stream.wait_event(slot["compute_done"])
                slot["x"].copy_(src, non_blocking=True)
                slot["h2d_done"].record(self.copy_stream)
            with torch.cuda.stream(self.stream):
                self.stream.wait_event(slot["h2d_done"])
                self.stream.wait_event(slot["d2h_done"])     # previous D2H of this slot's dets finished
                self._run_slot(slot)
                slot["compute_done"].record(self.stream)
            with torch.cuda.stream(self.d2h_stream):
                self.d2h_stream.wait_event(slot["compute_done"])
                slot["host_out"].copy_(slot["dets"], non_blocking=True)
                slot["d2h_done"].record(self.d2h_stream)
        self.copy_stream.synchronize()
        self.stream.synchronize()
        self.d2h_stream.synchronize()
        return self.slots[(n_steps - 1) % nslots]["host_out"]

    @property
    def h2d_bytes_per_step(self):
        b, c, h, w = self.shape
        return b * c * h * w * 4

    @property
    def d2h_bytes_per_step(self):
        return self.slots[0]["host_out"].numel() * 4
