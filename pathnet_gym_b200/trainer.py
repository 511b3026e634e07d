"""Population training step on one GPU's slice: forward + loss + backward for every local path,
sum of the shared gradient slab over ranks, one clipped RMSProp update.  This is the unit bench.py
times (one "step")."""
from . import parallel


class OverlappedGradSum(object):
    """Sum of the gradient slab over the GPUs, started early: the linear modules' region (97 % of the bytes) is
    complete before the conv layers' backward pass, so its all-reduce runs on a side stream underneath it; the two
    small regions (conv, heads) follow on the compute stream."""

    def __init__(self, pop):
        import os
        import torch
        self.pop = pop
        self.active = parallel.world()[1] > 1
        off, cnt = pop.lin_grads_region()
        g = pop.grads
        self.lin, self.head, self.conv = g[off:off + cnt], g[off + cnt:], g[:off]
        self.comm = torch.cuda.Stream(device=pop.device)
        self.comm_early = torch.cuda.Stream(device=pop.device)
        self.split = os.environ.get("PN_PEER_SPLIT", "0") == "1"      # PN_PEER_SPLIT=1: exchange all but the first layer early
        self.peer = None
        self.peer_slabs = None
        self.step = 0
        self.lin_step = 0
        if self.active:
            self._setup_peer_exchange()

    def _setup_peer_exchange(self):
        """Symmetric (peer-mapped) exchange buffers for the small regions; any failure leaves the NCCL path in place."""
        import os
        import torch
        import torch.distributed as dist
        if os.environ.get("PN_PEER_SUM", "1") == "0":     # PN_PEER_SUM=0: keep the grouped NCCL all-reduce
            return
        ok = torch.zeros(1, device=self.pop.device)
        try:
            import torch.distributed._symmetric_memory as symm
            nfloat = (self.pop.peer_buffer_bytes() + 3) // 4
            self._peer_buf = symm.empty(nfloat, dtype=torch.float32, device=self.pop.device)
            self._peer_buf.zero_()
            self._peer_hdl = symm.rendezvous(self._peer_buf, dist.group.WORLD)
            self.peer = [int(p) for p in self._peer_hdl.buffer_ptrs]
            if os.environ.get("PN_PEER_LIN", "0") == "1":
                # the gradient slab itself becomes symmetric memory: the linear region is then all-reduced in place
                # by pn_peer_allreduce_lin instead of NCCL
                g = symm.empty(self.pop.slab_floats, dtype=torch.float32, device=self.pop.device)
                g.copy_(self.pop.grads)
                self._slab_hdl = symm.rendezvous(g, dist.group.WORLD)
                self.peer_slabs = [int(p) for p in self._slab_hdl.buffer_ptrs]
                self.pop.rebind_grads(g)
                off, cnt = self.pop.lin_grads_region()
                self.lin, self.head, self.conv = g[off:off + cnt], g[off + cnt:], g[:off]
            ok.fill_(1)
        except Exception as e:      # noqa: BLE001  (old torch, no P2P, restricted container, ...)
            self.peer = None
            self.peer_slabs = None
            self.peer_error = repr(e)
        torch.cuda.synchronize(self.pop.device)
        dist.all_reduce(ok, op=dist.ReduceOp.MIN)       # all ranks take the same path; also orders the zero-fill
        if float(ok.item()) < 1:
            self.peer = None
            self.peer_slabs = None

    def check(self):
        """Synchronising health check of the peer-memory exchange (a rank that waited 2 s for a peer's flag gave up)."""
        if self.active and self.peer is not None:
            self.pop.peer_check(self.peer[parallel.world()[0]])

    def path_name(self):
        """Which exchange runs (for the bench line): nothing at one GPU; else the two regions' mechanisms."""
        if not self.active:
            return "none (1 GPU)"
        small = "NCCL all-reduce (grouped)"
        if self.peer is not None:
            small = "peer-memory kernel, middle conv + heads underneath the first layer's weight gradient, first layer " \
                    "after the step (pn_peer_sum_part)" if self.split else "peer-memory kernel (pn_peer_sum_small)"
        lin = "peer-memory kernel (pn_peer_allreduce_lin)" if (self.peer is not None and self.peer_slabs is not None) \
            else "NCCL all-reduce on a side stream"
        err = getattr(self, "peer_error", None)
        return "linear region: %s; conv + head regions: %s%s" % (lin, small, "; peer set-up failed: " + err if err else "")

    def sum_lin(self):
        """On the side stream (already waiting for the linear gradients): all-reduce them."""
        import os
        if self.active and "lin" not in os.environ.get("PN_DEBUG_SKIP_SUM", ""):     # (timing experiments only)
            if self.peer is not None and self.peer_slabs is not None:
                rank, world = parallel.world()
                self.pop.peer_allreduce_lin(self.peer_slabs, self.peer, rank, world, self.lin_step)
                self.lin_step += 1
                return
            import torch.distributed as dist
            dist.all_reduce(self.lin, op=dist.ReduceOp.SUM)

    def sum_rest(self):
        """On the compute stream, after the step's last kernel: one grouped NCCL launch for the two small regions."""
        import os
        if not self.active or "small" in os.environ.get("PN_DEBUG_SKIP_SUM", ""):    # (timing experiments only)
            return
        if self.peer is not None:
            rank, world = parallel.world()
            if self.split:
                import torch
                # everything but the first conv layer's tensors is final before that layer's weight-gradient kernel:
                # that part is exchanged on its own stream underneath it, only the rest after the step
                self.pop.stream_wait_early_grads(self.comm_early)
                with torch.cuda.stream(self.comm_early):
                    self.pop.peer_sum_part(self.peer, rank, world, self.step, 1)
                self.pop.peer_sum_part(self.peer, rank, world, self.step, 2)
                torch.cuda.current_stream(self.pop.device).wait_stream(self.comm_early)
            else:
                self.pop.peer_sum_small(self.peer, rank, world, self.step)
            self.step += 1
            return
        import torch.distributed as dist
        try:
            with dist._coalescing_manager(device=self.pop.device):
                dist.all_reduce(self.conv, op=dist.ReduceOp.SUM)
                dist.all_reduce(self.head, op=dist.ReduceOp.SUM)
        except (AttributeError, TypeError, RuntimeError):
            dist.all_reduce(self.conv, op=dist.ReduceOp.SUM)
            dist.all_reduce(self.head, op=dist.ReduceOp.SUM)

    def __call__(self):
        """Gradient sum only (no update): the slab holds the global sum when the compute stream gets there."""
        if not self.active:
            return
        import torch
        self.pop.stream_wait_lin_grads(self.comm)
        with torch.cuda.stream(self.comm):
            self.sum_lin()
        self.sum_rest()
        torch.cuda.current_stream(self.pop.device).wait_stream(self.comm)


class StepTail(object):
    """Everything between the backward pass and the next step: gradient sum over the GPUs + clipped RMSProp
    (a3c_training_thread.py:205-216 apply_gradients).  The linear modules' tensors -- gradients final right after their
    weight-gradient kernel -- are summed AND updated on a side stream while the conv layers' backward pass runs; the
    conv and head tensors follow on the compute stream.  Per-tensor results are the same as one apply_gradients call
    (tensors are independent in rmsprop_applier.py:92-106)."""

    def __init__(self, pop, decay=0.99, momentum=0.0, epsilon=0.1, clip_norm=40.0):
        self.pop = pop
        self.hp = (decay, momentum, epsilon, clip_norm)
        self.gradsum = OverlappedGradSum(pop)
        self.lin_first, self.lin_count = pop.lin_tensor_range()

    def __call__(self, lr):
        import torch
        pop, gs = self.pop, self.gradsum
        pop.stream_wait_lin_grads(gs.comm)
        with torch.cuda.stream(gs.comm):
            gs.sum_lin()
            pop.apply_gradients_range(self.lin_first, self.lin_count, lr, *self.hp)
        gs.sum_rest()
        pop.apply_gradients_range(self.lin_first, self.lin_count, lr, *self.hp, complement=True)
        torch.cuda.current_stream(pop.device).wait_stream(gs.comm)


class PopulationTrainer(object):
    def __init__(self, pop, entropy_beta=0.01, decay=0.99, momentum=0.0, epsilon=0.1, clip_norm=40.0):
        self.pop, self.beta = pop, entropy_beta
        self.decay, self.momentum, self.epsilon, self.clip_norm = decay, momentum, epsilon, clip_norm
        self.tail = StepTail(pop, decay, momentum, epsilon, clip_norm)

    def step(self, frames, actions, td, r, lr):
        """Device tensors [P,B,...]; returns the per-path losses (device tensor)."""
        loss = self.pop.train_step(frames, actions, td, r, self.beta)
        self.tail(lr)
        return loss

    def step_host(self, frames, actions, td, r, lr, loss_out=None):
        """Host (pinned) tensors: copies happen inside the C ABI call."""
        loss = self.pop.train_step_host(frames, actions, td, r, self.beta, loss_out=loss_out)
        self.tail(lr)
        return loss


class PipelinedHostTrainer(object):
    """Feeds host (pinned) batches through the double-buffered staging slots of the C ABI: while step i runs,
    batch i+1 crosses PCIe on a separate copy stream."""

    def __init__(self, pop, entropy_beta=0.01, decay=0.99, momentum=0.0, epsilon=0.1, clip_norm=40.0):
        import torch
        self.pop, self.beta = pop, entropy_beta
        self.decay, self.momentum, self.epsilon, self.clip_norm = decay, momentum, epsilon, clip_norm
        self.copy_stream = torch.cuda.Stream(device=pop.device)
        self.tail = StepTail(pop, decay, momentum, epsilon, clip_norm)
        self.i = 0
        self.primed = False

    def _stage(self, slot, batch):
        # a 4-D first tensor is a rollout's plane sequence [P, B + C - 1, H, W], a 5-D one is full stacks [P, B, H, W, C]
        if batch[0].dim() == 4:
            self.pop.stage_rollout_host(slot, *batch, copy_stream=self.copy_stream)
        else:
            self.pop.stage_host(slot, *batch, copy_stream=self.copy_stream)

    def prime(self, batch):
        """Stage the first batch (frames or planes, actions, td, r)."""
        self._stage(self.i & 1, batch)
        self.primed = True

    def step(self, next_batch, lr, loss_out=None):
        """Train on the batch staged last, and start staging `next_batch` (or None at the end)."""
        assert self.primed, "call prime() with the first batch"
        cur = self.i & 1
        if next_batch is not None:
            self._stage(cur ^ 1, next_batch)
        self.pop.train_step_staged(cur, self.beta, loss_out)
        self.tail(lr)
        self.i += 1
        self.primed = next_batch is not None
        return loss_out


class PopulationRollout(object):
    """The rollout loop of A3CTrainingThread.process (a3c_training_thread.py:118-180) for all paths of a population at
    once, with the bookkeeping on the device: one batched acting forward per environment step (run_policy_and_value),
    choose_action by `pn_sample_actions` (Philox stream instead of the global numpy RNG), reward clip + n-step returns +
    td by `pn_nstep_returns`.  Paths whose episode ends early stop collecting; their remaining slots are padding
    (action = -1).  `env_step(actions) -> (next_frames [P,H,W,C] u8 device tensor, rewards [P] float tensor,
    terminal [P] bool tensor)` is the caller's environment."""

    def __init__(self, pop, t_max, seed=0):
        import torch
        self.pop, self.t_max, self.seed = pop, int(t_max), int(seed)
        assert 1 <= self.t_max <= pop.B_max
        P, cfg, dev = pop.P, pop.cfg, pop.device
        self.frames = torch.zeros(P, self.t_max, cfg.H, cfg.W, cfg.C, dtype=torch.uint8, device=dev)
        self.actions = torch.full((P, self.t_max), -1, dtype=torch.int32, device=dev)
        self.values = torch.zeros(P, self.t_max, dtype=torch.float32, device=dev)
        self.rewards = torch.zeros(P, self.t_max, dtype=torch.float32, device=dev)
        self.epoch = 0

    def collect(self, state, env_step):
        """Runs up to t_max steps from `state` [P,H,W,C] u8; returns (frames, actions, td, R, lengths, last_state)."""
        import torch
        from .population import nstep_returns, sample_actions
        pop, P = self.pop, self.pop.P
        alive = torch.ones(P, dtype=torch.bool, device=pop.device)
        lengths = torch.zeros(P, dtype=torch.int32, device=pop.device)
        self.actions.fill_(-1)
        self.rewards.zero_()
        self.values.zero_()
        for i in range(self.t_max):
            pi, v = pop.forward(state.unsqueeze(1).contiguous())            # [P,1,A], [P,1]
            a = sample_actions(pi.reshape(P, -1).contiguous(), self.seed, self.epoch, id0=0)
            self.epoch += 1
            self.frames[:, i] = state
            self.actions[:, i] = torch.where(alive, a, torch.full_like(a, -1))
            self.values[:, i] = v.reshape(P)
            nxt, rew, term = env_step(a)
            self.rewards[:, i] = torch.where(alive, rew.to(torch.float32), torch.zeros_like(self.rewards[:, i]))
            lengths += alive.to(torch.int32)
            state = torch.where(alive.view(P, 1, 1, 1), nxt, state)
            alive = alive & ~term
            if not bool(alive.any()):
                break
        # bootstrap: V(s_T) for the paths still running, 0 after a terminal step (a3c_training_thread.py:151-153)
        _, vT = pop.forward(state.unsqueeze(1).contiguous())
        boot = torch.where(alive, vT.reshape(P), torch.zeros(P, device=pop.device)).contiguous()
        R, td = nstep_returns(self.rewards, self.values, boot, lengths)
        return self.frames, self.actions, td, R, lengths, state
