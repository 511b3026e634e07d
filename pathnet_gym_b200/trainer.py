"""Population training step on one GPU's slice: forward + loss + backward for every local path,
sum of the shared gradient slab over ranks, one clipped RMSProp update.  This is the unit bench.py
times (one "step")."""
from . import parallel


class OverlappedGradSum(object):
    """Sum of the gradient slab over the GPUs, started early: the linear modules' region (97 % of the bytes) is
    complete before the conv layers' backward pass, so its all-reduce runs on a side stream underneath it; the two
    small regions (conv, heads) follow on the compute stream."""

    def __init__(self, pop):
        import torch
        self.pop = pop
        self.active = parallel.world()[1] > 1
        if self.active:
            off, cnt = pop.lin_grads_region()
            g = pop.grads
            self.lin, self.head, self.conv = g[off:off + cnt], g[off + cnt:], g[:off]
            self.comm = torch.cuda.Stream(device=pop.device)

    def __call__(self):
        if not self.active:
            return
        import torch
        import torch.distributed as dist
        self.pop.stream_wait_lin_grads(self.comm)
        with torch.cuda.stream(self.comm):
            dist.all_reduce(self.lin, op=dist.ReduceOp.SUM)
        # the two small regions are complete only when the step's last kernel is: one grouped NCCL launch for both
        try:
            with dist._coalescing_manager(device=self.pop.device):
                dist.all_reduce(self.conv, op=dist.ReduceOp.SUM)
                dist.all_reduce(self.head, op=dist.ReduceOp.SUM)
        except (AttributeError, TypeError, RuntimeError):
            dist.all_reduce(self.conv, op=dist.ReduceOp.SUM)
            dist.all_reduce(self.head, op=dist.ReduceOp.SUM)
        torch.cuda.current_stream(self.pop.device).wait_stream(self.comm)


class PopulationTrainer(object):
    def __init__(self, pop, entropy_beta=0.01, decay=0.99, momentum=0.0, epsilon=0.1, clip_norm=40.0):
        self.pop, self.beta = pop, entropy_beta
        self.decay, self.momentum, self.epsilon, self.clip_norm = decay, momentum, epsilon, clip_norm
        self.gradsum = OverlappedGradSum(pop)

    def step(self, frames, actions, td, r, lr):
        """Device tensors [P,B,...]; returns the per-path losses (device tensor)."""
        loss = self.pop.train_step(frames, actions, td, r, self.beta)
        self.gradsum()
        self.pop.apply_gradients(lr, self.decay, self.momentum, self.epsilon, self.clip_norm)
        return loss

    def step_host(self, frames, actions, td, r, lr, loss_out=None):
        """Host (pinned) tensors: copies happen inside the C ABI call."""
        loss = self.pop.train_step_host(frames, actions, td, r, self.beta, loss_out=loss_out)
        self.gradsum()
        self.pop.apply_gradients(lr, self.decay, self.momentum, self.epsilon, self.clip_norm)
        return loss


class PipelinedHostTrainer(object):
    """Feeds host (pinned) batches through the double-buffered staging slots of the C ABI: while step i runs,
    batch i+1 crosses PCIe on a separate copy stream."""

    def __init__(self, pop, entropy_beta=0.01, decay=0.99, momentum=0.0, epsilon=0.1, clip_norm=40.0):
        import torch
        self.pop, self.beta = pop, entropy_beta
        self.decay, self.momentum, self.epsilon, self.clip_norm = decay, momentum, epsilon, clip_norm
        self.copy_stream = torch.cuda.Stream(device=pop.device)
        self.gradsum = OverlappedGradSum(pop)
        self.i = 0
        self.primed = False

    def prime(self, batch):
        """Stage the first batch (frames, actions, td, r)."""
        self.pop.stage_host(self.i & 1, *batch, copy_stream=self.copy_stream)
        self.primed = True

    def step(self, next_batch, lr, loss_out=None):
        """Train on the batch staged last, and start staging `next_batch` (or None at the end)."""
        assert self.primed, "call prime() with the first batch"
        cur = self.i & 1
        if next_batch is not None:
            self.pop.stage_host(cur ^ 1, *next_batch, copy_stream=self.copy_stream)
        self.pop.train_step_staged(cur, self.beta, loss_out)
        self.gradsum()
        self.pop.apply_gradients(lr, self.decay, self.momentum, self.epsilon, self.clip_norm)
        self.i += 1
        self.primed = next_batch is not None
        return loss_out
