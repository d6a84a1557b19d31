"""Learning-rate policy of the DLRM training loop: linear warm-up, constant plateau, quadratic decay.

Mirror of ``LRPolicyScheduler`` (reference ``P/dlrm_s_pytorch.py:133-167``) with the same constructor
arguments, so ``lr_scheduler.get_last_lr()[0]`` can feed ``emb_l.sgd_args._replace(lr=...)``
(``P/dlrm_s_pytorch.py:1027``) unchanged.  One deliberate difference: with ``num_warmup_steps == 0`` and a
decay that starts later, the reference reads ``self.last_lr`` before ever setting it (AttributeError);
here the plateau simply holds the base rate.
"""
from __future__ import annotations

import sys

from torch.optim.lr_scheduler import _LRScheduler


class LRPolicyScheduler(_LRScheduler):
    def __init__(self, optimizer, num_warmup_steps, decay_start_step, num_decay_steps):
        self.num_warmup_steps = num_warmup_steps
        self.decay_start_step = decay_start_step
        self.decay_end_step = decay_start_step + num_decay_steps
        self.num_decay_steps = num_decay_steps
        if self.decay_start_step < self.num_warmup_steps:
            sys.exit("Learning rate warmup must finish before the decay starts")
        self._frozen = None
        super().__init__(optimizer)

    def get_lr(self):
        step = self._step_count
        if step < self.num_warmup_steps:
            scale = 1.0 - (self.num_warmup_steps - step) / self.num_warmup_steps
            lr = [b * scale for b in self.base_lrs]
            self._frozen = lr
        elif self.decay_start_step <= step < self.decay_end_step:
            scale = ((self.num_decay_steps - (step - self.decay_start_step)) / self.num_decay_steps) ** 2
            lr = [max(0.0000001, b * scale) for b in self.base_lrs]
            self._frozen = lr
        elif self.num_decay_steps > 0 and self._frozen is not None:
            lr = self._frozen  # between warm-up and decay, or after the decay: hold the last value
        else:
            lr = list(self.base_lrs)
        return lr
