"""Adam with the update done by the fused xnn_adam_step_multi kernel (one launch per parameter group) (K14: architectures/__init__.py:463-464).

State layout is torch.optim.Adam's ({"step", "exp_avg", "exp_avg_sq"} per parameter), so optimizer
state_dicts inside reference checkpoints (selene/__init__.py:376-384,208-221) load and save unchanged.

Device-resident step (`enable_device_step`): the bias corrections are computed IN the kernel from an int64 counter in
device memory that a one-thread kernel advances, so a captured CUDA graph of the training step does not bake the
host's step count into its launch parameters.  The host-side "step" entries keep counting (`note_replays` after graph
replays) so that `state_dict()` stays what torch.optim.Adam would have written.
"""
import torch

from . import ops


class FusedAdam(torch.optim.Optimizer):
    def __init__(self, params, lr=1e-3, betas=(0.9, 0.999), eps=1e-8, weight_decay=0.0):
        super().__init__(params, dict(lr=lr, betas=betas, eps=eps, weight_decay=weight_decay))
        self._step_dev = None

    def _host_step(self):
        steps = [int(st["step"]) for st in self.state.values() if "step" in st]
        return max(steps) if steps else 0

    def enable_device_step(self, device):
        """All parameters then share ONE device-resident step counter (they are stepped together, as the
        reference's optimizer.step() does)."""
        if self._step_dev is None or self._step_dev.device != torch.device(device):
            self._step_dev = torch.full((1,), self._host_step(), dtype=torch.int64, device=device)
        return self._step_dev

    def note_replays(self, n=1):
        """n CUDA-graph replays of a captured step() happened: advance the host-side step entries"""
        for st in self.state.values():
            if "step" in st:
                st["step"] += n

    def load_state_dict(self, state_dict):
        super().load_state_dict(state_dict)
        if self._step_dev is not None:
            self._step_dev.fill_(self._host_step())

    @torch.no_grad()
    def step(self, closure=None, grad_scale=1.0):
        loss = None
        if closure is not None:
            with torch.enable_grad():
                loss = closure()
        if self._step_dev is not None:
            ops.counters_advance(self._step_dev)
        for group in self.param_groups:
            b1, b2 = group["betas"]
            by_step = {}            # parameters stepped together share one launch
            for p in group["params"]:
                if p.grad is None:
                    continue
                st = self.state[p]
                if len(st) == 0:
                    st["step"] = torch.tensor(0.0)
                    st["exp_avg"] = torch.zeros_like(p, memory_format=torch.contiguous_format)
                    st["exp_avg_sq"] = torch.zeros_like(p, memory_format=torch.contiguous_format)
                st["step"] += 1
                g = p.grad if p.grad.is_contiguous() else p.grad.contiguous()
                key = 0 if self._step_dev is not None else int(st["step"])
                lst = by_step.setdefault(key, ([], [], [], []))
                lst[0].append(p.data); lst[1].append(g); lst[2].append(st["exp_avg"]); lst[3].append(st["exp_avg_sq"])
            for step, (ps, gs, ms, vs) in by_step.items():
                ops.adam_step_multi(ps, gs, ms, vs, group["lr"], b1, b2, group["eps"], group["weight_decay"], step,
                                    grad_scale, step_dev=self._step_dev)
        return loss
