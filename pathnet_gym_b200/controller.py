"""Population controller: the evolution + task-transfer semantics of the reference's controller worker
(doom_pathnet.py:202-296) without its parameter-server polling.

State is replicated on every rank (scores [P_total] f32, genotypes [P_total, L] u32 bitmasks, fixed path
[L]); every rank runs the same deterministic pn_tournament (same seed / epoch) and then loads its own
slice of the effective masks into its PathNetPopulation."""
import numpy as np
import torch

from . import parallel
from . import population as _pop

SCORE_UNSET = -1000.0     # doom_pathnet.py:131,267


class PopulationController(object):
    def __init__(self, pop, P_total, Bcand=3, seed=1, rank=0, world_size=1):
        self.pop, self.cfg = pop, pop.cfg
        self.P_total, self.Bcand, self.seed = int(P_total), int(Bcand), int(seed)
        self.rank, self.world_size = rank, world_size
        self.lo, self.hi = parallel.shard_range(P_total, rank, world_size)
        assert self.hi - self.lo == pop.P, "population slice does not match the handle's P"
        dev = pop.device
        self.epoch = 0
        self.fixed = torch.zeros(self.cfg.L, dtype=torch.int32, device=dev)
        self.fixed_path = np.zeros((self.cfg.L, self.cfg.M))
        self.scores = torch.full((self.P_total,), SCORE_UNSET, dtype=torch.float32, device=dev)
        self.genotypes = None
        self.winner_idx = 0
        self.vars_backup = pop.params.clone()          # vars_backup = sess.run(vars_) (doom_pathnet.py:204-205)

    def _next_epoch(self):
        self.epoch += 1
        return self.epoch

    def _load_slice(self, effective):
        self.pop.set_genotypes(effective[self.lo:self.hi].cpu().numpy().view(np.uint32))

    def start_task(self):
        """Fresh random genotypes OR the fixed path for every trainer (doom_pathnet.py:212-221)."""
        cfg = self.cfg
        self.genotypes = _pop.geopath_init(self.P_total, cfg.L, cfg.M, cfg.N, self.seed, self._next_epoch(),
                                           device=self.pop.device_index)
        self.scores.fill_(SCORE_UNSET)
        self._load_slice(self.genotypes | self.fixed[None, :])

    def report_scores(self, local_scores):
        """Each rank contributes the episode scores of its slice (score_set_ops, a3c_training_thread.py:143-147)."""
        sizes = [parallel.shard_range(self.P_total, r, self.world_size) for r in range(self.world_size)]
        allsc = parallel.allgather_concat(local_scores.to(self.scores.device), [b - a for a, b in sizes])
        self.scores.copy_(allsc)

    def tournament_step(self, n_groups=None):
        """One controller pass (doom_pathnet.py:241-270).  n_groups=1 is one reference tournament; the
        default runs P_total // Bcand disjoint tournaments at once.  Returns the winners (numpy, -1 = group
        skipped because a candidate had not reported a score)."""
        cfg = self.cfg
        if n_groups is None:
            n_groups = self.P_total // self.Bcand
        eff, winners = _pop.tournament_(self.genotypes, self.scores, self.fixed, cfg.M, cfg.N, self.Bcand, n_groups,
                                        self.seed, self._next_epoch())
        self._load_slice(eff)
        w = winners.cpu().numpy()
        ok = w[w >= 0]
        if ok.size:
            self.winner_idx = int(ok[-1])       # last completed tournament, as the reference's winner_idx
        return w

    def end_task(self):
        """Task switch (doom_pathnet.py:273-293): fixed_path <- winner's genotype (replaces, no OR);
        its tensors are frozen; every other tensor (heads included) is re-initialised from the backup;
        RMSProp slots are kept."""
        cfg = self.cfg
        win = self.genotypes[self.winner_idx].clone()
        self.fixed = win
        m = win.cpu().numpy().view(np.uint32)
        self.fixed_path = ((m[:, None] >> np.arange(cfg.M, dtype=np.uint32)[None, :]) & 1).astype(float)
        self.pop.set_fixed_path(self.fixed_path)
        self.pop.reinit_unfrozen(self.vars_backup)
        return self.fixed_path
