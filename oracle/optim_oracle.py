"""TEST INFRASTRUCTURE ONLY (tests/, smoke, bench CPU legs): CPU statement of the reference's optimizer step.

The reference's optimizer IS torch's: ``torch.optim.AdamW(self.parameters(), lr=1e-4, weight_decay=0)`` (FiT/train.py:215-217), followed by
the EMA wrapper's update ``_foreach_mul_(ema, decay); _foreach_add_(ema, params, alpha=1 - decay)`` (FiT/ema.py:149-154), run when
``current_step % every_n_steps == 0`` (ema.py:252-257).  So the oracle calls torch.optim.AdamW on CPU fp32 (pinned by construction: it is
the same third-party code the reference executes) and restates the two EMA lines."""
import torch


class AdamWEMAOracle:
    def __init__(self, params, lr=1e-4, betas=(0.9, 0.999), eps=1e-8, weight_decay=0.0, ema_decay=0.9999, every_n_steps=1):
        self.params = [torch.nn.Parameter(p.detach().clone().float().cpu()) for p in params]
        self.opt = torch.optim.AdamW(self.params, lr=lr, betas=betas, eps=eps, weight_decay=weight_decay, foreach=False)
        self.ema = [p.detach().clone() for p in self.params] if ema_decay is not None else []
        self.decay, self.every, self.current_step = ema_decay, every_n_steps, 0

    @torch.no_grad()
    def step(self, grads):
        for p, g in zip(self.params, grads):
            p.grad = g.detach().clone().float().cpu()
        self.opt.step()
        if self.ema and self.current_step % self.every == 0:   # ema.py:252-257 then ema_update (ema.py:149-154)
            torch._foreach_mul_(self.ema, self.decay)
            torch._foreach_add_(self.ema, [p.detach() for p in self.params], alpha=1.0 - self.decay)
        self.current_step += 1

    def state(self, i):
        st = self.opt.state[self.params[i]]
        return st["exp_avg"], st["exp_avg_sq"]
