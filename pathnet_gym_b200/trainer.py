"""Population training step on one GPU's slice: forward + loss + backward for every local path,
sum of the shared gradient slab over ranks, one clipped RMSProp update.  This is the unit bench.py
times (one "step")."""
from . import parallel


class PopulationTrainer(object):
    def __init__(self, pop, entropy_beta=0.01, decay=0.99, momentum=0.0, epsilon=0.1, clip_norm=40.0):
        self.pop, self.beta = pop, entropy_beta
        self.decay, self.momentum, self.epsilon, self.clip_norm = decay, momentum, epsilon, clip_norm

    def step(self, frames, actions, td, r, lr):
        """Device tensors [P,B,...]; returns the per-path losses (device tensor)."""
        loss = self.pop.train_step(frames, actions, td, r, self.beta)
        parallel.allreduce_sum_(self.pop.grads)
        self.pop.apply_gradients(lr, self.decay, self.momentum, self.epsilon, self.clip_norm)
        return loss

    def step_host(self, frames, actions, td, r, lr, loss_out=None):
        """Host (pinned) tensors: copies happen inside the C ABI call."""
        loss = self.pop.train_step_host(frames, actions, td, r, self.beta, loss_out=loss_out)
        parallel.allreduce_sum_(self.pop.grads)
        self.pop.apply_gradients(lr, self.decay, self.momentum, self.epsilon, self.clip_norm)
        return loss
