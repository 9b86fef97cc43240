"""Tokenizer service path: latents in pinned host memory -> codes and quantized latents in pinned host memory.

`quantizer.encode(x)` / `.decode(codes)` work on device tensors, like the reference.  A caller whose latents
arrive in host memory (a data loader, an RPC front end) pays the PCIe copies around them; done one after
the other they cost several times the kernels (49 MB in, 52 MB out per MusicGen batch of 64 x 30 s).
`HostPipeline` splits the batch into chunks of whole items and runs three CUDA streams -- host->device copy,
encode+decode, device->host copy -- so that the copies of neighbouring chunks (and of neighbouring calls)
overlap each other and the kernels; PCIe is full duplex.  Frames are independent, so chunking does not change
a single code.
"""
from __future__ import annotations

import typing as tp

import torch


class HostPipeline:
    def __init__(self, quantizer, device: tp.Optional[torch.device] = None, chunk_items: int = 32, slots: int = 3):
        self.q = quantizer
        self.device = torch.device(device) if device is not None else next(quantizer.buffers()).device
        if self.device.type != "cuda":
            raise RuntimeError("HostPipeline needs the quantizer on a CUDA device (no CPU fallback)")
        self.chunk_items = int(chunk_items)
        self.slots = int(slots)
        self.s_in = torch.cuda.Stream(self.device)
        self.s_cmp = torch.cuda.Stream(self.device)
        self.s_out = torch.cuda.Stream(self.device)
        self._x_dev: tp.List[torch.Tensor] = []
        self._free: tp.List[tp.Optional[torch.cuda.Event]] = []  # slot may be overwritten after this event
        self._next_slot = 0

    def _staging(self, D: int, T: int) -> None:
        """Staging slots hold `chunk_items * D * T` floats and only ever grow: a call with shorter items uses views of
        the same memory.  When they do have to grow, the copies and kernels that still use the old slots are waited
        for first (the old blocks go back to the allocator, which may hand them out again at once)."""
        need = self.chunk_items * D * T
        if self._x_dev and self._x_dev[0].numel() >= need:
            return
        for ev in self._free:
            if ev is not None:
                ev.synchronize()
        self.s_in.synchronize()
        self._x_dev = [torch.empty(need, dtype=torch.float32, device=self.device) for _ in range(self.slots)]
        self._free = [None] * self.slots
        self._next_slot = 0

    @torch.no_grad()
    def encode_decode(self, x_host: torch.Tensor, codes_host: torch.Tensor, quant_host: torch.Tensor) -> torch.cuda.Event:
        """x_host [B, D, T] fp32 -> codes_host [B, K, T] int64 and quant_host [B, T, D] fp32 (the physical layout
        of the reference's [B, D, T] result), all pinned host tensors.  Asynchronous: returns the event that
        completes when the last byte is in host memory (`event.synchronize()` or `self.synchronize()`)."""
        for t, name in ((x_host, "x_host"), (codes_host, "codes_host"), (quant_host, "quant_host")):
            if t.is_cuda or not t.is_pinned() or not t.is_contiguous():
                raise ValueError(f"{name} must be a contiguous pinned host tensor")
        B, D, T = x_host.shape
        self._staging(D, T)
        # the codes decoded here come straight from encode(), so the host-side range check (one stream
        # synchronisation per decode) is switched off for the duration of the call
        vq = getattr(self.q, "vq", None)
        saved = getattr(vq, "check_codes", None)
        if saved is not None:
            vq.check_codes = False
        cur = torch.cuda.current_stream(self.device)
        self.s_in.wait_stream(cur)  # whatever produced the host data on the caller's stream is done first
        last = None
        try:
            last = self._run_chunks(x_host, codes_host, quant_host, B, D, T)
        finally:
            if saved is not None:
                vq.check_codes = saved
        return last

    def _run_chunks(self, x_host, codes_host, quant_host, B, D, T):
        last = None
        for lo in range(0, B, self.chunk_items):
            hi = min(B, lo + self.chunk_items)
            slot = self._next_slot
            self._next_slot = (slot + 1) % self.slots
            xd = self._x_dev[slot][: (hi - lo) * D * T].view(hi - lo, D, T)
            with torch.cuda.stream(self.s_in):
                if self._free[slot] is not None:
                    self.s_in.wait_event(self._free[slot])  # the kernels that read this staging slot last
                xd.copy_(x_host[lo:hi], non_blocking=True)
                ev_in = torch.cuda.Event()
                ev_in.record(self.s_in)
            with torch.cuda.stream(self.s_cmp):
                self.s_cmp.wait_event(ev_in)
                codes = self.q.encode(xd)                       # [b, K, T] view of a [K, b, T] block
                quant = self.q.decode(codes).transpose(1, 2)    # contiguous [b, T, D]
                codes_c = codes.contiguous()
                ev_cmp = torch.cuda.Event()
                ev_cmp.record(self.s_cmp)
                self._free[slot] = ev_cmp
            with torch.cuda.stream(self.s_out):
                self.s_out.wait_event(ev_cmp)
                codes_host[lo:hi].copy_(codes_c, non_blocking=True)
                quant_host[lo:hi].copy_(quant, non_blocking=True)
                # the device results were allocated on s_cmp: keep the allocator from reusing them too early
                codes_c.record_stream(self.s_out)
                quant.record_stream(self.s_out)
                last = torch.cuda.Event(enable_timing=True)
                last.record(self.s_out)
        return last

    def synchronize(self) -> None:
        self.s_out.synchronize()
