"""Adam with the update done by the fused xnn_adam_step_multi kernel (one launch per parameter group) (K14: architectures/__init__.py:463-464).

State layout is torch.optim.Adam's ({"step", "exp_avg", "exp_avg_sq"} per parameter), so optimizer
state_dicts inside reference checkpoints (selene/__init__.py:376-384,208-221) load and save unchanged.
"""
import torch

from . import ops


class FusedAdam(torch.optim.Optimizer):
    def __init__(self, params, lr=1e-3, betas=(0.9, 0.999), eps=1e-8, weight_decay=0.0):
        super().__init__(params, dict(lr=lr, betas=betas, eps=eps, weight_decay=weight_decay))

    @torch.no_grad()
    def step(self, closure=None, grad_scale=1.0):
        loss = None
        if closure is not None:
            with torch.enable_grad():
                loss = closure()
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
                lst = by_step.setdefault(int(st["step"]), ([], [], [], []))
                lst[0].append(p.data); lst[1].append(g); lst[2].append(st["exp_avg"]); lst[3].append(st["exp_avg_sq"])
            for step, (ps, gs, ms, vs) in by_step.items():
                ops.adam_step_multi(ps, gs, ms, vs, group["lr"], b1, b2, group["eps"], group["weight_decay"], step,
                                    grad_scale)
        return loss
