"""Flat trainable state: one contiguous fp32 buffer for params, one for grads, two for Adam.

The reference keeps 16 separate trainable tensors and lets DDP bucket their grads
(4,659,528 B in two buckets) or all-reduces 116 tensors one by one
(X6).  On B200 the aggregation kernels want exactly one peer-mapped buffer per
client, so the trainable parameters are re-pointed as views into a single flat
allocation (optionally living in symmetric memory), and the gradient views are
pre-installed as ``param.grad`` so autograd accumulates straight into the flat
gradient buffer.  Segment offsets are padded to 128 B so every tensor starts on a
vectorisable boundary.
"""
from __future__ import annotations

from typing import Callable, Dict, List, Optional, Tuple

import torch
from torch import nn

ALIGN_ELEMS = 32  # 128 bytes of fp32


class FlatState:
    def __init__(self, named_params: List[Tuple[str, nn.Parameter]], device: torch.device,
                 allocator: Optional[Callable[[int], torch.Tensor]] = None, with_adam: bool = True,
                 with_grad: bool = True):
        self.names = [n for n, _ in named_params]
        self.params = [p for _, p in named_params]
        self.offsets: Dict[str, Tuple[int, int]] = {}
        off = 0
        for n, p in named_params:
            self.offsets[n] = (off, p.numel())
            off += (p.numel() + ALIGN_ELEMS - 1) // ALIGN_ELEMS * ALIGN_ELEMS
        self.numel = off
        self.payload_numel = sum(p.numel() for p in self.params)
        alloc = allocator or (lambda n: torch.zeros(n, dtype=torch.float32, device=device))
        self.param = alloc(self.numel)
        self.param.zero_()
        self.grad = None
        if with_grad:
            self.grad = alloc(self.numel)
            self.grad.zero_()
        if with_adam:
            self.exp_avg = torch.zeros(self.numel, dtype=torch.float32, device=device)
            self.exp_avg_sq = torch.zeros(self.numel, dtype=torch.float32, device=device)
        self.step = 0
        with torch.no_grad():
            for n, p in named_params:
                o, k = self.offsets[n]
                view = self.param[o:o + k].view(p.shape)
                view.copy_(p.data.to(device))
                p.data = view
                if with_grad:
                    p.grad = self.grad[o:o + k].view(p.shape)

    def view(self, name: str, grad: bool = False) -> torch.Tensor:
        o, k = self.offsets[name]
        buf = self.grad if grad else self.param
        p = self.params[self.names.index(name)]
        return buf[o:o + k].view(p.shape)

    def rebind(self) -> None:
        """Re-install the views (after ``load_state_dict`` or ``zero_grad(set_to_none=True)``)."""
        for n, p in zip(self.names, self.params):
            o, k = self.offsets[n]
            if p.data.data_ptr() != self.param[o:o + k].data_ptr():
                with torch.no_grad():
                    self.param[o:o + k].view(p.shape).copy_(p.data)
                p.data = self.param[o:o + k].view(p.shape)
            if self.grad is not None:
                p.grad = self.grad[o:o + k].view(p.shape)

    def zero_grad(self) -> None:
        self.grad.zero_()

    def grad_sq_norm(self) -> torch.Tensor:
        return (self.grad * self.grad).sum()
