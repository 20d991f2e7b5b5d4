"""Training-step tail on the GPU: gradient-norm clipping + Adam in two
kernel launches over all parameter tensors.

Drop-in for the optimizer the reference builds at train.py:129-131

    optimizier = torch.optim.Adam(model.parameters(), lr=args.lr, weight_decay=args.l2)

and for the clip call of solver.py:184-185.  With max_norm given, step() performs
torch.nn.utils.clip_grad_norm_(params, max_norm) and the Adam update together; a caller that
still clips beforehand (as solver.py does) gets the same result, because already-clipped
gradients have norm <= max_norm and the coefficient is 1.  state_dict() / load_state_dict() use
torch.optim.Adam's format (step, exp_avg, exp_avg_sq per parameter; param_groups with lr, betas,
eps, weight_decay), so checkpoints written by the reference load and vice versa; solver.py:116-119
halves the learning rate through that dict.
"""
import ctypes as C
import math

import numpy as np
import torch

from . import _native as nat

_CHUNK = 8192   # elements per CTA
# struct ctn_opt_chunk as a numpy record (include/ctn_b200.h)
_CHUNK_DTYPE = np.dtype([("param", "<u8"), ("grad", "<u8"), ("exp_avg", "<u8"), ("exp_avg_sq", "<u8"),
                         ("n", "<i4"), ("pad_", "<i4")])


class FusedAdam(torch.optim.Optimizer):
    """Adam (L2 weight decay, amsgrad off) with optional fused global-norm gradient clipping.
    CUDA float32 parameters only; one parameter group."""

    def __init__(self, params, lr=1e-3, betas=(0.9, 0.999), eps=1e-8, weight_decay=0.0, max_norm=None,
                 write_clipped_grad=False):
        if lr < 0.0 or eps < 0.0 or not (0.0 <= betas[0] < 1.0) or not (0.0 <= betas[1] < 1.0) or weight_decay < 0.0:
            raise ValueError("Invalid Adam hyper-parameter")   # torch.optim.Adam raises ValueError too
        defaults = dict(lr=lr, betas=tuple(betas), eps=eps, weight_decay=weight_decay, amsgrad=False, maximize=False,
                        foreach=None, capturable=False, differentiable=False, fused=None)
        super().__init__(params, defaults)
        if len(self.param_groups) != 1:
            raise ValueError("FusedAdam supports one parameter group")
        self.max_norm = max_norm
        self.write_clipped_grad = bool(write_clipped_grad)
        self.total_norm = None      # device scalar: gradient norm before clipping of the last step
        self._steps = 0
        self._flat_m = self._flat_v = None
        self._table = self._key_p = self._key_g = self._partials = None

    # ---------------------------------------------------------------- state
    def _init_state(self):
        ps = self.param_groups[0]["params"]
        for p in ps:
            if not p.is_cuda or p.dtype != torch.float32 or not p.is_contiguous():
                raise RuntimeError("FusedAdam needs contiguous CUDA float32 parameters")
        dev = ps[0].device
        n = sum(p.numel() for p in ps)
        self._flat_m = torch.zeros(n, dtype=torch.float32, device=dev)
        self._flat_v = torch.zeros(n, dtype=torch.float32, device=dev)
        self.total_norm = torch.zeros((), dtype=torch.float32, device=dev)
        off = 0
        self._step_t = torch.tensor(float(self._steps))   # one host tensor shared by every parameter's state
        for p in ps:
            k = p.numel()
            self.state[p] = {"step": self._step_t,
                             "exp_avg": self._flat_m[off:off + k].view_as(p),
                             "exp_avg_sq": self._flat_v[off:off + k].view_as(p)}
            off += k

    def state_dict(self):
        """torch.optim.Adam's format.  Every parameter gets its OWN step tensor: torch's Adam increments
        each parameter's step in place, so a shared one would advance len(params) times per step."""
        sd = super().state_dict()
        for s in sd["state"].values():
            if "step" in s:
                s["step"] = torch.tensor(float(self._steps))
        return sd

    def load_state_dict(self, state_dict):
        """Accepts torch.optim.Adam's format; the moments are copied into the flat buffers."""
        super().load_state_dict(state_dict)
        loaded = {p: dict(s) for p, s in self.state.items()}
        self._flat_m = None
        self.state.clear()
        self._init_state()
        steps = 0
        for p, s in loaded.items():
            if "exp_avg" in s:
                self.state[p]["exp_avg"].copy_(s["exp_avg"])
                self.state[p]["exp_avg_sq"].copy_(s["exp_avg_sq"])
                steps = max(steps, int(float(s.get("step", 0))))
        self._steps = steps
        self._step_t.fill_(float(steps))
        self._table = None

    # ---------------------------------------------------------------- step
    def _build_table(self, ps):
        chunks = []
        for p in ps:
            if p.grad.dtype != torch.float32 or not p.grad.is_contiguous() or p.grad.device != p.device:
                raise RuntimeError("FusedAdam needs contiguous float32 gradients on the parameter's device")
            st = self.state[p]
            n = p.numel()
            for lo in range(0, n, _CHUNK):
                c = nat.OptChunk()
                c.param = p.data_ptr() + 4 * lo
                c.grad = p.grad.data_ptr() + 4 * lo
                c.exp_avg = st["exp_avg"].data_ptr() + 4 * lo
                c.exp_avg_sq = st["exp_avg_sq"].data_ptr() + 4 * lo
                c.n = min(_CHUNK, n - lo)
                chunks.append(c)
        arr = (nat.OptChunk * len(chunks))(*chunks)
        dev = ps[0].device
        host = np.frombuffer(bytearray(bytes(arr)), dtype=_CHUNK_DTYPE)
        self._table = torch.from_numpy(host.view(np.uint8)).to(dev)
        # the grad column as 64-bit integers on the device: when only the gradients move (and all by the
        # same offset: they are views of one flat buffer that the backward allocates anew every step) the
        # column is shifted in place by one asynchronous launch, instead of rebuilding the table chunk by
        # chunk (3.5 ms of host time) or re-uploading it from pageable memory (waits for the stream)
        words = _CHUNK_DTYPE.itemsize // 8
        self._grad_col = self._table.view(torch.int64).view(-1, words)[:, _CHUNK_DTYPE.fields["grad"][1] // 8]
        self._n_chunks = len(chunks)
        self._partials = torch.empty(len(chunks), dtype=torch.float64, device=dev)

    def _refresh_table(self, ps):
        key_p = tuple(p.data_ptr() for p in ps)
        key_g = tuple(p.grad.data_ptr() for p in ps)
        if self._table is not None and key_p == self._key_p:
            if key_g == self._key_g:
                return
            delta = key_g[0] - self._key_g[0]
            if all(a - b == delta for a, b in zip(key_g, self._key_g)) and all(
                    p.grad.dtype == torch.float32 and p.grad.is_contiguous() for p in ps):
                self._grad_col.add_(delta)   # pointer arithmetic only; wraps like the addresses do
                self._key_g = key_g
                return
        self._build_table(ps)
        self._key_p, self._key_g = key_p, key_g

    @torch.no_grad()
    def step(self, closure=None):
        loss = None
        if closure is not None:
            with torch.enable_grad():
                loss = closure()
        g = self.param_groups[0]
        if self._flat_m is None:
            self._init_state()
        ps = [p for p in g["params"] if p.grad is not None]
        if not ps:
            return loss
        self._refresh_table(ps)
        self._steps += 1
        t = self._steps
        b1, b2 = g["betas"]
        step_size = g["lr"] / (1.0 - b1 ** t)
        inv_sqrt_bias2 = 1.0 / math.sqrt(1.0 - b2 ** t)
        with torch.cuda.device(ps[0].device):
            nat.check(nat.lib().ctn_clip_adam(
                self._table.data_ptr(), self._n_chunks, self._partials.data_ptr(),
                float(self.max_norm) if self.max_norm is not None else 0.0, step_size, b1, b2, inv_sqrt_bias2,
                g["eps"], g["weight_decay"], int(self.write_clipped_grad), self.total_norm.data_ptr(), nat.stream()))
        # the kernel wrote the parameters through raw device pointers: bump their version counters, which is
        # what the engines' prepared (bf16 split / norm-folded / transposed) weight copies are keyed on
        # (_engine._Prepared.refresh), exactly as an in-place torch update would have
        torch.autograd.graph.increment_version(ps)
        self._step_t += 1      # host tensor, as torch.optim.Adam keeps it by default
        return loss
