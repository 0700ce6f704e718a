"""Flat parameter / gradient arenas and the fused Adam step (train.py:95,161; train_avse_frames.py:91,180).

B200 layout: every trainable tensor of a model is a view into ONE contiguous fp32 buffer (16-byte aligned slices),
and so is every ``.grad``.  Consequences: the optimizer is one streaming kernel over the arena (28 B/param), the
data-parallel gradient exchange is one NCCL all-reduce over the gradient arena (no bucketing copies), and
``zero_grad`` is one memset.  ``Adam`` keeps torch.optim.Adam's constructor and ``state_dict`` layout
(``exp_avg`` / ``exp_avg_sq`` / ``step`` per parameter, as views of the arenas) so reference checkpoints
(utilities.py:171-204) load.
"""
import torch

from ._lib import check, dptr, fptr, lib, stream_ptr

_ALIGN = 4   # floats (16 bytes)


class FlatParams:
    """Re-point the given parameters (unique, in order) at slices of one arena and give each a .grad view."""

    def __init__(self, params):
        seen, uniq = set(), []
        for p in params:
            if id(p) not in seen:
                seen.add(id(p))
                uniq.append(p)
        if not uniq:
            raise ValueError("no parameters")
        dev = uniq[0].device
        if dev.type != "cuda":
            raise RuntimeError("FlatParams needs CUDA parameters (maavss_b200 has no CPU path)")
        offs, n = [], 0
        for p in uniq:
            if p.dtype != torch.float32 or p.device != dev:
                raise RuntimeError("all parameters must be fp32 on one device")
            offs.append(n)
            n += (p.numel() + _ALIGN - 1) // _ALIGN * _ALIGN
        self.params, self.offsets, self.numel = uniq, offs, n
        self.index = {id(p): i for i, p in enumerate(uniq)}
        self.data = torch.zeros(n, device=dev, dtype=torch.float32)
        self.grad = torch.zeros(n, device=dev, dtype=torch.float32)
        with torch.no_grad():
            for p, o in zip(uniq, offs):
                v = self.data[o:o + p.numel()].view(p.shape)
                v.copy_(p.data)
                p.data = v
                p.grad = self.grad[o:o + p.numel()].view(p.shape)

    def view_like(self, arena, i):
        p, o = self.params[i], self.offsets[i]
        return arena[o:o + p.numel()].view(p.shape)

    def zero_grad(self):
        check(lib().maavss_fill(fptr(self.grad), self.numel, 0.0, stream_ptr()))
        for i, p in enumerate(self.params):          # re-attach views a caller may have dropped (set_to_none)
            if p.grad is None or p.grad.data_ptr() != self.grad.data_ptr() + 4 * self.offsets[i]:
                p.grad = self.view_like(self.grad, i)


class Adam(torch.optim.Optimizer):
    """torch.optim.Adam(params, lr, betas, eps, weight_decay) semantics, one kernel per step over the flat arena.

    Parameters whose ``requires_grad`` is False at step time keep their value (their arena gradient is zero and so
    are their moments unless they were trained before: Adam's update of a zero gradient with zero moments is zero)."""

    def __init__(self, params, lr=1e-3, betas=(0.9, 0.999), eps=1e-8, weight_decay=0.0, grad_scale=1.0):
        params = list(params)
        if params and isinstance(params[0], dict):
            raise NotImplementedError("parameter groups are not supported: one flat arena, one set of hyper-parameters")
        super().__init__(params, dict(lr=lr, betas=betas, eps=eps, weight_decay=weight_decay))
        self.flat = FlatParams(self.param_groups[0]["params"])
        self.exp_avg = torch.zeros_like(self.flat.data)
        self.exp_avg_sq = torch.zeros_like(self.flat.data)
        self.step_count = 0          # host mirror of the device counter
        self.dev_state = torch.zeros(2, device=self.flat.data.device, dtype=torch.int64)   # {step; step_size, 1/sqrt(bc2)}
        self.grad_scale = grad_scale
        for i, p in enumerate(self.flat.params):
            self.state[p] = {"step": torch.tensor(0.0), "exp_avg": self.flat.view_like(self.exp_avg, i),
                             "exp_avg_sq": self.flat.view_like(self.exp_avg_sq, i)}

    def zero_grad(self, set_to_none=False):
        self.flat.zero_grad()

    @torch.no_grad()
    def prepare(self):
        """First half of a step: advance the device step counter / bias corrections (once per optimizer step)."""
        g = self.param_groups[0]
        self.step_count += 1
        check(lib().maavss_adam_prepare(dptr(self.dev_state), g["lr"], g["betas"][0], g["betas"][1], stream_ptr()))

    @torch.no_grad()
    def apply_range(self, lo=0, hi=None, zero_grad=False, max_ctas=0):
        """Second half: update arena elements [lo, hi) (multiples of 4) with the prepared step; optionally clear their gradients."""
        hi = self.flat.numel if hi is None else hi
        if hi <= lo:
            return
        if lo % _ALIGN:
            raise ValueError("apply_range: lo must be a multiple of 4 elements")
        g = self.param_groups[0]
        f = self.flat
        check(lib().maavss_adam_apply(fptr(f.data[lo:hi]), fptr(f.grad[lo:hi]), fptr(self.exp_avg[lo:hi]), fptr(self.exp_avg_sq[lo:hi]),
                                      hi - lo, g["betas"][0], g["betas"][1], g["eps"], g["weight_decay"], dptr(self.dev_state),
                                      self.grad_scale, int(zero_grad), max_ctas, stream_ptr()))

    @torch.no_grad()
    def step(self, closure=None):
        loss = closure() if closure is not None else None
        self.prepare()
        self.apply_range()
        return loss

    @torch.no_grad()
    def step_and_zero_grad(self):
        """optimizer.step(); optimizer.zero_grad() (train.py:161-162) as one pass over the arenas."""
        self.prepare()
        self.apply_range(zero_grad=True)

    def _sync_step_state(self):
        """The device counter is the truth (a captured train step advances it on every replay without passing through
        step()); the per-parameter `step` entries of torch.optim.Adam's state layout are filled from it on demand."""
        self.step_count = int(self.dev_state[0].item())
        for st in self.state.values():
            st["step"] = torch.tensor(float(self.step_count))

    def state_dict(self):
        self._sync_step_state()
        return super().state_dict()

    def load_state_dict(self, state_dict):
        """Accepts a torch.optim.Adam state_dict: copies the moments into the arenas (keeps the views)."""
        ids = state_dict["param_groups"][0]["params"]
        for idx, p in zip(ids, self.param_groups[0]["params"]):
            st = state_dict["state"].get(idx)
            if st is None:
                continue
            i = self.flat.index[id(p)]
            self.flat.view_like(self.exp_avg, i).copy_(st["exp_avg"])
            self.flat.view_like(self.exp_avg_sq, i).copy_(st["exp_avg_sq"])
            self.step_count = max(self.step_count, int(float(st["step"])))
        self.dev_state[0] = self.step_count
        g, src = self.param_groups[0], state_dict["param_groups"][0]
        for k in ("lr", "betas", "eps", "weight_decay"):
            if k in src:
                g[k] = src[k]
