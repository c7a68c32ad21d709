"""Engine: one resident model per (process, GPU), driven through the C ABI.

Replaces the per-task ``ResNet50(weights=...)`` / ``InceptionV3(weights=...)`` +
``predict`` of /root/reference/models.py:25-26, :40, :50-51, :65.  PyTorch is used
for device/pinned allocations and the CUDA stream only; every kernel is ours and
runs through libdml_b200.so.  No CPU fallback: without a GPU this raises.
"""
from __future__ import annotations

import ctypes as C
import threading

import numpy as np

from . import _lib, frontend, keras_weights


def _torch():
    import torch
    return torch


class Engine:
    def __init__(self, model_name: str, device: int = 0, max_batch: int = 256, micro_batch: int = 0,
                 fp32_check: bool = False, tensor_cores: bool = True, keep_all: bool = False,
                 weights=None, load: bool = True):
        torch = _torch()
        if not torch.cuda.is_available():
            raise _lib.DmlError("no CUDA device: the dml_b200 engine has no CPU fallback")
        self.l = _lib.lib()
        self.lock = threading.Lock()  # one engine is driven by one thread at a time
        self.model_name = model_name
        self.device = device
        self.max_batch = max_batch
        self.size, self.mode = _lib.model_info(model_name)
        flags = (_lib.FLAG_FP32_CHECK if fp32_check else 0) | (0 if tensor_cores else _lib.FLAG_NO_TENSOR_CORE) | \
                (_lib.FLAG_KEEP_ALL if keep_all else 0)
        self.fp32 = fp32_check
        self.h = C.c_void_p()
        torch.cuda.set_device(device)
        _lib.check(self.l.dml_create(device, _lib.model_id(model_name), max_batch, micro_batch, flags, C.byref(self.h)))
        ws, wt = C.c_size_t(), C.c_size_t()
        _lib.check(self.l.dml_workspace_bytes(self.h, C.byref(ws)))
        _lib.check(self.l.dml_weights_bytes(self.h, C.byref(wt)))
        dev = torch.device("cuda", device)
        self.stream = torch.cuda.Stream(device=dev)
        self.workspace = torch.empty(ws.value + 256, dtype=torch.uint8, device=dev)
        self.weights = torch.empty(wt.value + 256, dtype=torch.uint8, device=dev)
        self._ws_ptr = (self.workspace.data_ptr() + 255) // 256 * 256
        self._wt_ptr = (self.weights.data_ptr() + 255) // 256 * 256
        self.weights_bytes = wt.value
        _lib.check(self.l.dml_bind_weights(self.h, self._wt_ptr, wt.value))
        _lib.check(self.l.dml_bind_workspace(self.h, self._ws_ptr, ws.value))
        # device-side inputs/outputs, sized for max_batch
        S = self.size
        self.max_pixels = max_batch * 300 * 300 * 3
        # a single image above this is resized on the host through the same index tables (frontend.shrink_oversized)
        self.max_image_bytes = max(S * S * 3, self.max_pixels // min(max_batch, 8))
        self._host_slots = None
        self.d_pixels = torch.empty(self.max_pixels + 16, dtype=torch.uint8, device=dev)
        self.d_off = torch.empty(max_batch, dtype=torch.int64, device=dev)
        self.d_hw = torch.empty((max_batch, 2), dtype=torch.int32, device=dev)
        self.d_xtab = torch.empty((max_batch, S), dtype=torch.int32, device=dev)
        self.d_ytab = torch.empty((max_batch, S), dtype=torch.int32, device=dev)
        # second slot of inputs/results + a copy stream: H2D of batch i+1 overlaps the kernels of batch i
        self.copy_stream = torch.cuda.Stream(device=dev)
        self._slots = None
        self._submitted = 0
        self.d_probs = torch.empty((max_batch, 1000), dtype=torch.float32, device=dev)
        self.d_idx = torch.empty((max_batch, 5), dtype=torch.int32, device=dev)
        self.d_prob = torch.empty((max_batch, 5), dtype=torch.float32, device=dev)
        self.h_idx = torch.empty((max_batch, 5), dtype=torch.int32).pin_memory()
        self.h_prob = torch.empty((max_batch, 5), dtype=torch.float32).pin_memory()
        if load:
            self.load_weights(weights)

    # -- weights ---------------------------------------------------------
    def load_weights(self, weights=None):
        """weights: flat fp32 blob, list of arrays in blob order, or None (file/env or fixed-seed default)."""
        if weights is None:
            blob = keras_weights.load_blob(self.model_name)
        elif isinstance(weights, np.ndarray) and weights.ndim == 1:
            blob = np.ascontiguousarray(weights, dtype=np.float32)
        else:
            blob = keras_weights.to_blob(weights)
        _lib.check(self.l.dml_load_weights(self.h, blob.ctypes.data, blob.size))

    def packed_weights(self):
        """The packed device weights as a torch uint8 view (for an NCCL broadcast)."""
        off = self._wt_ptr - self.weights.data_ptr()
        return self.weights[off:off + self.weights_bytes]

    def weights_ready(self):
        _lib.check(self.l.dml_weights_ready(self.h))

    # -- hot path --------------------------------------------------------
    def make_current(self):
        """CUDA's current device is per THREAD: a task that runs in an executor thread starts on device 0, and the first
        allocation or event there would build a context on GPU 0 from every worker process (1.7 s, once, per process --
        measured on the 8-GPU job mix).  Every entry point that may be the first CUDA call of its thread selects the
        engine's device first."""
        _torch().cuda.set_device(self.device)

    def upload(self, hb: frontend.HostBatch):
        """Host->device copy of one HostBatch on the engine stream."""
        torch = _torch()
        self.make_current()
        n = hb.n
        if n > self.max_batch:
            raise _lib.DmlError(f"batch of {n} exceeds max_batch {self.max_batch}")
        if hb.pixels.size > self.d_pixels.numel():
            raise _lib.DmlError("pixel blob larger than the device staging buffer")
        with torch.cuda.stream(self.stream):
            self.d_pixels[:hb.pixels.size].copy_(torch.from_numpy(hb.pixels), non_blocking=True)
            self.d_off[:n].copy_(torch.from_numpy(hb.img_off), non_blocking=True)
            self.d_hw[:n].copy_(torch.from_numpy(hb.hw), non_blocking=True)
            self.d_xtab[:n].copy_(torch.from_numpy(hb.xtab), non_blocking=True)
            self.d_ytab[:n].copy_(torch.from_numpy(hb.ytab), non_blocking=True)

    def preprocess(self, n: int):
        _lib.check(self.l.dml_preprocess(self.h, self.d_pixels.data_ptr(), self.d_off.data_ptr(), self.d_hw.data_ptr(),
                                         self.d_xtab.data_ptr(), self.d_ytab.data_ptr(), n, self.stream.cuda_stream))

    def forward(self, n: int, want_probs: bool = True):
        _lib.check(self.l.dml_forward(self.h, n, self.d_probs.data_ptr() if want_probs else None,
                                      self.d_idx.data_ptr(), self.d_prob.data_ptr(), self.stream.cuda_stream))

    def infer_device(self, n: int, want_probs: bool = False):
        """dml_infer on what upload() put on the device; results stay on the device."""
        _lib.check(self.l.dml_infer(self.h, self.d_pixels.data_ptr(), self.d_off.data_ptr(), self.d_hw.data_ptr(),
                                    self.d_xtab.data_ptr(), self.d_ytab.data_ptr(), n,
                                    self.d_probs.data_ptr() if want_probs else None, self.d_idx.data_ptr(),
                                    self.d_prob.data_ptr(), self.stream.cuda_stream))

    def infer_ptrs(self, pixels, img_off, hw, xtab, ytab, n: int, want_probs: bool = False):
        """dml_infer over caller-owned device tensors (uint8 pixels, int64 offsets, int32 sizes/tables)."""
        _lib.check(self.l.dml_infer(self.h, pixels.data_ptr(), img_off.data_ptr(), hw.data_ptr(), xtab.data_ptr(),
                                    ytab.data_ptr(), n, self.d_probs.data_ptr() if want_probs else None,
                                    self.d_idx.data_ptr(), self.d_prob.data_ptr(), self.stream.cuda_stream))

    def pin(self, hb: frontend.HostBatch) -> frontend.HostBatch:
        """Moves a HostBatch's arrays into pinned host memory (async H2D copies need it)."""
        torch = _torch()
        for name in ("pixels", "img_off", "hw", "xtab", "ytab"):
            t = torch.from_numpy(getattr(hb, name)).pin_memory()
            setattr(hb, name, t.numpy())
            setattr(hb, "_pin_" + name, t)
        return hb

    def predict_host_batch(self, hb: frontend.HostBatch):
        """One step end to end: H2D of the batch, dml_infer, D2H of the top-5, sync.
        Returns views of the pinned result buffers (valid until the next call)."""
        torch = _torch()
        self.upload(hb)
        self.infer_device(hb.n)
        with torch.cuda.stream(self.stream):
            self.h_idx[:hb.n].copy_(self.d_idx[:hb.n], non_blocking=True)
            self.h_prob[:hb.n].copy_(self.d_prob[:hb.n], non_blocking=True)
        self.stream.synchronize()
        return self.h_idx[:hb.n].numpy(), self.h_prob[:hb.n].numpy()

    # -- pipelined host-buffer path -------------------------------------------
    def _make_slots(self):
        torch = _torch()
        dev = self.d_pixels.device
        S, mb = self.size, self.max_batch
        slots = []
        for k in range(2):
            sl = {
                "pixels": self.d_pixels if k == 0 else torch.empty_like(self.d_pixels),
                "img_off": self.d_off if k == 0 else torch.empty_like(self.d_off),
                "hw": self.d_hw if k == 0 else torch.empty_like(self.d_hw),
                "xtab": self.d_xtab if k == 0 else torch.empty_like(self.d_xtab),
                "ytab": self.d_ytab if k == 0 else torch.empty_like(self.d_ytab),
                "d_idx": torch.empty((mb, 5), dtype=torch.int32, device=dev),
                "d_prob": torch.empty((mb, 5), dtype=torch.float32, device=dev),
                "h_idx": torch.empty((mb, 5), dtype=torch.int32).pin_memory(),
                "h_prob": torch.empty((mb, 5), dtype=torch.float32).pin_memory(),
                "copied": torch.cuda.Event(), "done": torch.cuda.Event(), "n": 0, "busy": False,
            }
            slots.append(sl)
        self._slots = slots

    def submit(self, hb: frontend.HostBatch) -> int:
        """Enqueue one batch (host buffers, ideally pinned): H2D on the copy stream, dml_infer and the
        D2H of the top-5 on the compute stream.  Returns a ticket for collect().  At most two batches
        may be in flight; the H2D of a batch overlaps the kernels of the one before it."""
        torch = _torch()
        self.make_current()
        if self._slots is None:
            self._make_slots()
        k = self._submitted % 2
        sl = self._slots[k]
        if sl["busy"]:
            raise _lib.DmlError("two batches already in flight: collect() one first")
        n = hb.n
        if n > self.max_batch or hb.pixels.size > sl["pixels"].numel():
            raise _lib.DmlError("batch larger than the engine's staging buffers")
        with torch.cuda.stream(self.copy_stream):
            self.copy_stream.wait_event(sl["done"])  # the slot's previous kernels have consumed its inputs
            sl["pixels"][:hb.pixels.size].copy_(torch.from_numpy(hb.pixels), non_blocking=True)
            sl["img_off"][:n].copy_(torch.from_numpy(hb.img_off), non_blocking=True)
            sl["hw"][:n].copy_(torch.from_numpy(hb.hw), non_blocking=True)
            sl["xtab"][:n].copy_(torch.from_numpy(hb.xtab), non_blocking=True)
            sl["ytab"][:n].copy_(torch.from_numpy(hb.ytab), non_blocking=True)
            sl["copied"].record(self.copy_stream)
        with torch.cuda.stream(self.stream):
            self.stream.wait_event(sl["copied"])
            _lib.check(self.l.dml_infer(self.h, sl["pixels"].data_ptr(), sl["img_off"].data_ptr(), sl["hw"].data_ptr(),
                                        sl["xtab"].data_ptr(), sl["ytab"].data_ptr(), n, None, sl["d_idx"].data_ptr(),
                                        sl["d_prob"].data_ptr(), self.stream.cuda_stream))
            sl["h_idx"][:n].copy_(sl["d_idx"][:n], non_blocking=True)
            sl["h_prob"][:n].copy_(sl["d_prob"][:n], non_blocking=True)
            sl["done"].record(self.stream)
        sl["n"], sl["busy"] = n, True
        self._submitted += 1
        return k

    def collect(self, ticket: int):
        """Wait for a submitted batch; returns (top5_idx, top5_prob) views of pinned memory, valid until the
        slot is reused two submits later."""
        sl = self._slots[ticket]
        if not sl["busy"]:
            raise _lib.DmlError("nothing in flight for this ticket")
        sl["done"].synchronize()
        sl["busy"] = False
        n = sl["n"]
        return sl["h_idx"][:n].numpy(), sl["h_prob"][:n].numpy()

    def chunks(self, images):
        """Splits a list of uint8 images into runs that fit one engine batch: at most max_batch images AND at most
        the staging buffer's bytes (full-resolution pixels are uploaded, so the byte bound matters for anything larger
        than the 300x300 pictures the reference bundles).  Oversized single images are shrunk on the host first."""
        run, nbytes = [], 0
        for im in images:
            im = frontend.shrink_oversized(im, self.size, self.max_image_bytes)
            if run and (len(run) == self.max_batch or nbytes + im.nbytes > self.max_pixels):
                yield run
                run, nbytes = [], 0
            run.append(im)
            nbytes += im.nbytes
        if run:
            yield run

    def host_slot(self, k: int) -> dict:
        """Pinned host arrays of slot k (two slots): frontend.HostBatch(..., into=slot) fills them in place, so
        the H2D copies of submit() are asynchronous without a pin_memory() call per batch."""
        torch = _torch()
        self.make_current()
        if self._host_slots is None:
            S, mb = self.size, self.max_batch
            self._host_slots = []
            for _ in range(2):
                t = {"pixels": torch.empty(self.max_pixels + 16, dtype=torch.uint8).pin_memory(),
                     "img_off": torch.empty(mb, dtype=torch.int64).pin_memory(),
                     "hw": torch.empty((mb, 2), dtype=torch.int32).pin_memory(),
                     "xtab": torch.empty((mb, S), dtype=torch.int32).pin_memory(),
                     "ytab": torch.empty((mb, S), dtype=torch.int32).pin_memory()}
                self._host_slots.append(({k2: v.numpy() for k2, v in t.items()}, t))
        return self._host_slots[k][0]

    def predict_pipelined(self, images, cancelled=None):
        """uint8 HxWx3 arrays (any iterable, consumed lazily) -> (top5_idx [n,5], top5_prob [n,5]) through
        submit()/collect(): while the GPU runs batch i, the caller's iterator (JPEG decoding) and the host-side
        packing of batch i+1 proceed, and its H2D copy overlaps the kernels.  ``cancelled``: an optional
        threading.Event checked between batches (the worker's task preemption, worker.py:944-953)."""
        idx_out, prob_out = [], []
        pending = None

        def drain():
            nonlocal pending
            if pending is not None:
                i, p = self.collect(pending)
                idx_out.append(i.copy())
                prob_out.append(p.copy())
                pending = None

        try:
            for run in self.chunks(images):
                if cancelled is not None and cancelled.is_set():
                    raise _lib.DmlCancelled("cancelled between two batches")
                hb = frontend.HostBatch(run, self.size, into=self.host_slot(self._submitted % 2))
                t = self.submit(hb)
                drain()
                pending = t
            drain()
        except BaseException:
            if pending is not None:  # leave no batch in flight behind an error
                try:
                    self.collect(pending)
                except Exception:
                    pass
            raise
        if not idx_out:
            return np.zeros((0, 5), np.int32), np.zeros((0, 5), np.float32)
        return np.concatenate(idx_out), np.concatenate(prob_out)

    def cancel(self):
        """Stops the dml_infer call in progress (or the next one) before its next micro-batch."""
        _lib.check(self.l.dml_cancel(self.h))

    def predict_images(self, images, want_probs: bool = False):
        """uint8 HxWx3 arrays -> (top5_idx [n,5] int32, top5_prob [n,5] fp32[, probs [n,1000]])."""
        torch = _torch()
        idx_out, prob_out, probs_out = [], [], []
        for chunk in self.chunks(images):
            hb = frontend.HostBatch(chunk, self.size)
            self.upload(hb)
            self.infer_device(hb.n, want_probs)
            with torch.cuda.stream(self.stream):
                self.h_idx[:hb.n].copy_(self.d_idx[:hb.n], non_blocking=True)
                self.h_prob[:hb.n].copy_(self.d_prob[:hb.n], non_blocking=True)
            self.stream.synchronize()
            idx_out.append(self.h_idx[:hb.n].numpy().copy())
            prob_out.append(self.h_prob[:hb.n].numpy().copy())
            if want_probs:
                probs_out.append(self.d_probs[:hb.n].cpu().numpy())
        if not idx_out:
            e = (np.zeros((0, 5), np.int32), np.zeros((0, 5), np.float32))
            return e + ((np.zeros((0, 1000), np.float32),) if want_probs else ())
        r = (np.concatenate(idx_out), np.concatenate(prob_out))
        return r + ((np.concatenate(probs_out),) if want_probs else ())

    # -- inspection ------------------------------------------------------
    def input_tensor(self, n: int):
        torch = _torch()
        p, eb = C.c_void_p(), C.c_int()
        _lib.check(self.l.dml_input_tensor(self.h, C.byref(p), C.byref(eb)))
        off = p.value - self.workspace.data_ptr()
        layout, cells, cch = C.c_int(), C.c_int(), C.c_int()
        _lib.check(self.l.dml_input_layout(self.h, C.byref(layout), C.byref(cells), C.byref(cch)))
        if layout.value == 1:
            # space-to-depth cells [n][S2][row_cells][(py, px, c) + 4 zero channels] -> the [n][S][S][3] picture; rows are padded
            # (data cell x at column first + x) and cells whose physical column has bit `bit` set store their halves swapped
            s2 = cells.value
            rc, first, bit = C.c_int(), C.c_int(), C.c_int()
            _lib.check(self.l.dml_input_rows(self.h, C.byref(rc), C.byref(first), C.byref(bit)))
            rows = self.workspace[off:off + n * s2 * rc.value * 16 * 2].view(torch.bfloat16).view(n, s2, rc.value, 16)
            cols = torch.arange(first.value, first.value + s2, device=rows.device)
            raw = rows[:, :, first.value:first.value + s2, :]
            if bit.value >= 0:
                swapped = ((cols >> bit.value) & 1).bool().view(1, 1, s2, 1)
                raw = torch.where(swapped, torch.cat([raw[..., 8:], raw[..., :8]], dim=-1), raw)
            pic = raw[..., :12].reshape(n, s2, s2, 2, 2, 3).permute(0, 1, 3, 2, 4, 5).reshape(n, 2 * s2, 2 * s2, 3)
            return pic[:, :self.size, :self.size].contiguous()
        nbytes = n * self.size * self.size * 3 * eb.value
        raw = self.workspace[off:off + nbytes]
        dt = torch.float32 if eb.value == 4 else torch.bfloat16
        return raw.view(dt).view(n, self.size, self.size, 3)

    def logits(self, n: int):
        torch = _torch()
        p = C.c_void_p()
        _lib.check(self.l.dml_logits(self.h, C.byref(p)))
        off = p.value - self.workspace.data_ptr()
        return self.workspace[off:off + n * 4000].view(torch.float32).view(n, 1000)

    def tensors(self):
        """{name: (index, h, w, c, c_pitch, elem_bytes, dev_ptr)}"""
        n = C.c_int()
        _lib.check(self.l.dml_num_tensors(self.h, C.byref(n)))
        out = {}
        buf = C.create_string_buffer(128)
        for i in range(n.value):
            h, w, c, cp, eb, keep = (C.c_int() for _ in range(6))
            p = C.c_void_p()
            _lib.check(self.l.dml_tensor_info(self.h, i, buf, 128, C.byref(h), C.byref(w), C.byref(c), C.byref(p),
                                              C.byref(cp), C.byref(eb), C.byref(keep)))
            out[buf.value.decode()] = (i, h.value, w.value, c.value, cp.value, eb.value, p.value)
        return out

    def read_tensor(self, name: str, n: int):
        """Copy a tensor of the current micro-batch to the host as fp32 NHWC (keep_all engines)."""
        torch = _torch()
        _, h, w, c, cp, eb, ptr = self.tensors()[name]
        off = ptr - self.workspace.data_ptr()
        nbytes = ((n * h * w - 1) * cp + c) * eb
        raw = self.workspace[off:off + nbytes + (-nbytes) % eb]
        dt = torch.float32 if eb == 4 else torch.bfloat16
        flat = raw.view(dt)
        full = torch.zeros(n * h * w * cp, dtype=dt, device=flat.device)
        full[:flat.numel()] = flat
        return full.view(n, h, w, cp)[..., :c].float().cpu().numpy()

    def ops(self):
        """[(name, kind, impl, macs_per_img, act_bytes_per_img, weight_bytes)] of the layer plan."""
        n = C.c_int()
        _lib.check(self.l.dml_num_ops(self.h, C.byref(n)))
        buf = C.create_string_buffer(128)
        out = []
        for i in range(n.value):
            k, im = C.c_int(), C.c_int()
            macs, ab, wb = C.c_int64(), C.c_int64(), C.c_int64()
            _lib.check(self.l.dml_op_info(self.h, i, buf, 128, C.byref(k), C.byref(im), C.byref(macs), C.byref(ab),
                                          C.byref(wb)))
            out.append((buf.value.decode(), k.value, im.value, macs.value, ab.value, wb.value))
        return out

    def schedule(self):
        """[(name, stream, num_waits)] per layer-plan op and the number of streams in use (dml_op_schedule)."""
        ops = self.ops()
        out, ns = [], C.c_int(1)
        for i, o in enumerate(ops[:-2]):  # the dense and softmax ops always run on the caller's stream
            st, nw = C.c_int(), C.c_int()
            _lib.check(self.l.dml_op_schedule(self.h, i, C.byref(st), C.byref(nw), C.byref(ns)))
            out.append((o[0], st.value, nw.value))
        return out, ns.value

    def profile_forward(self, n: int):
        """Per-op milliseconds of one forward over the current input tensor (CUDA events around each launch)."""
        nops = len(self.ops())
        ms = np.zeros(nops, dtype=np.float32)
        _lib.check(self.l.dml_profile_forward(self.h, n, self.d_idx.data_ptr(), self.d_prob.data_ptr(), ms.ctypes.data,
                                              nops, self.stream.cuda_stream))
        return ms

    def launch_count(self) -> int:
        v = C.c_int64()
        _lib.check(self.l.dml_launch_count(self.h, C.byref(v)))
        return v.value

    def sync(self):
        self.stream.synchronize()

    def close(self):
        if getattr(self, "h", None) and self.h.value:
            self.sync()
            self.l.dml_destroy(self.h)
            self.h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass
