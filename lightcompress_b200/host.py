"""Fake-quant of weights that live in HOST memory (checkpoint shards, CPU-offloaded blocks): the call a user makes when
the model does not sit on the GPU — LLMC keeps blocks on the CPU between `block_opt` calls (blockwise_optimization.py:
53-61, base_bq.py:397-418) and quantizes one block at a time.

Row chunks of every tensor flow  pinned host -> H2D (stream 1) -> fused quant-dequant kernel (stream 2) -> D2H (stream 3)
-> pinned host  through a small ring of device buffers, so that the PCIe link is busy in both directions at once and the
pipeline's fill / drain cost is one chunk, not one tensor. The arithmetic is `fake_quant_weight_dynamic` per chunk
(quantization units are rows / groups inside a row, so row chunks are independent).
"""
import torch

from . import _cabi


class HostFakeQuant(object):
    def __init__(self, quantizer, device, chunk_bytes=16 << 20, depth=4):
        if quantizer.granularity not in ('per_group', 'per_channel', 'per_token'):
            raise NotImplementedError('HostFakeQuant: units must lie inside one row (per_group / per_channel / per_token)')
        if quantizer.calib_algo != 'minmax':
            # the mse / learnable paths allocate their own output (quant.py `_dynamic_mse`): the ring buffers would be
            # copied back unwritten
            raise NotImplementedError(f"HostFakeQuant: calib_algo={quantizer.calib_algo!r} (only 'minmax' writes in place)")
        self.q = quantizer
        self.dev = torch.device(device)
        self.chunk_bytes, self.depth = int(chunk_bytes), int(depth)
        self.s_in, self.s_k, self.s_out = (torch.cuda.Stream(self.dev) for _ in range(3))
        self.din = [torch.empty(self.chunk_bytes, dtype=torch.uint8, device=self.dev) for _ in range(self.depth)]
        self.dout = [torch.empty(self.chunk_bytes, dtype=torch.uint8, device=self.dev) for _ in range(self.depth)]
        self.ev_in = [torch.cuda.Event() for _ in range(self.depth)]
        self.ev_k = [torch.cuda.Event() for _ in range(self.depth)]
        self.ev_out = [torch.cuda.Event() for _ in range(self.depth)]
        self.k = 0
        self.launches = 0

    def run(self, hin, hout):
        """hin / hout: lists of pinned host tensors [rows, cols] (same shapes / dtype). Asynchronous: call synchronize()."""
        for src, dst in zip(hin, hout):
            if not (src.is_pinned() and dst.is_pinned()):
                raise ValueError('HostFakeQuant: host tensors must be pinned')
            rows, cols = src.shape
            row_bytes = cols * src.element_size()
            step = max(1, self.chunk_bytes // row_bytes)
            if row_bytes > self.chunk_bytes:
                raise ValueError('HostFakeQuant: chunk_bytes is smaller than one row')
            for r0 in range(0, rows, step):
                r1 = min(rows, r0 + step)
                n = (r1 - r0) * row_bytes
                slot = self.k % self.depth
                first = self.k < self.depth
                a = self.din[slot][:n].view(src.dtype).view(r1 - r0, cols)
                b = self.dout[slot][:n].view(src.dtype).view(r1 - r0, cols)
                with torch.cuda.stream(self.s_in):
                    if not first:
                        self.s_in.wait_event(self.ev_k[slot])      # the kernel that read this input buffer is done
                    a.copy_(src[r0:r1], non_blocking=True)
                    self.ev_in[slot].record(self.s_in)
                with torch.cuda.stream(self.s_k):
                    self.s_k.wait_event(self.ev_in[slot])
                    if not first:
                        self.s_k.wait_event(self.ev_out[slot])     # the D2H that read this output buffer is done
                    self.q._dynamic(a, _cabi.OUT_FAKE, a.dtype, False, out=b)
                    self.ev_k[slot].record(self.s_k)
                with torch.cuda.stream(self.s_out):
                    self.s_out.wait_event(self.ev_k[slot])
                    dst[r0:r1].copy_(b, non_blocking=True)
                    self.ev_out[slot].record(self.s_out)
                self.k += 1
                self.launches += 1

    def begin(self, stream=None):
        """Order the pipeline after everything already queued on `stream` (default: the current stream)."""
        stream = stream or torch.cuda.current_stream(self.dev)
        ev = torch.cuda.Event()
        ev.record(stream)
        for s in (self.s_in, self.s_k, self.s_out):
            s.wait_event(ev)

    def end(self, stream=None):
        """Make `stream` (default: the current stream) wait for every chunk queued so far (results are in host memory)."""
        stream = stream or torch.cuda.current_stream(self.dev)
        for ev in self.ev_out[:min(self.k, self.depth)]:
            stream.wait_event(ev)

    def synchronize(self):
        for s in (self.s_in, self.s_k, self.s_out):
            s.synchronize()
