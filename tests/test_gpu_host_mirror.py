"""GPU tests of the host-side mirror of the reference interface (pathnet, game_ac_network,
a3c_training_thread, rmsprop_applier, controller) -- they read like the reference's call sites."""
import types

import numpy as np
import pytest

torch = pytest.importorskip("torch")

from oracle import pathnet_oracle as po
from tests import helpers as hp

pytestmark = pytest.mark.gpu


def make_flags(**kw):
    d = dict(task_index=0, L=3, M=10, N=3, kernel_num="8,4,3", stride_size="4,2,1", worker_hosts_num=4, H=84, W=84,
             population=1, batch_max=20, B=3)
    d.update(kw)
    return types.SimpleNamespace(**d)


def test_pathnet_module_matches_oracle(native_lib):
    from pathnet_gym_b200 import pathnet
    pathnet.seed(1234, epoch=5)
    g = pathnet.get_geopath(4, 10, 4)
    assert g.shape == (4, 10) and np.all(g.sum(axis=1) == 4)
    assert np.array_equal(pathnet.masks_from_geopath(g), po.get_geopath_masks(4, 10, 4, 1234, 5, 0))
    g2 = pathnet.mutation(g.copy(), 4, 10, 4)
    assert np.array_equal(pathnet.masks_from_geopath(g2), po.mutate_masks(pathnet.masks_from_geopath(g), 4, 10, 4, 1234, 6, 0))
    assert np.array_equal(pathnet.geopath_initializer(3, 4), np.ones((3, 4)))


def test_network_thread_applier_like_the_reference_driver(native_lib):
    """doom_pathnet.py:95-113,186-199: build applier + training thread, set a genotype, run process()."""
    from pathnet_gym_b200 import pathnet
    from pathnet_gym_b200.a3c_training_thread import A3CTrainingThread
    from pathnet_gym_b200.constants import GRAD_NORM_CLIP, MAX_TIME_STEP, RMSP_ALPHA, RMSP_EPSILON
    from pathnet_gym_b200.rmsprop_applier import RMSPropApplier
    FLAGS = make_flags(use_lstm=False)            # the feed-forward class (the reference constant picks the LSTM class)
    np.random.seed(0)
    applier = RMSPropApplier(learning_rate=None, decay=RMSP_ALPHA, momentum=0.0, epsilon=RMSP_EPSILON,
                             clip_norm=GRAD_NORM_CLIP, device="/gpu:0")
    th = A3CTrainingThread(0, "", 0, 7.0016e-4, None, applier, MAX_TIME_STEP, device="/gpu:0", FLAGS=FLAGS, task_index=0)
    net = th.local_network
    assert len(net.get_vars()) == 2 * (2 * 10 + 10) + 4 == len(th.apply_gradients)
    pathnet.seed(7)
    geo = pathnet.get_geopath(FLAGS.L, FLAGS.M, FLAGS.N)
    pathnet.geopath_insert(None, net.geopath_update_placeholders_set[0], net.geopath_update_ops_set[0], geo, FLAGS.L, FLAGS.M)
    assert np.array_equal(net.get_geopath(None), geo)
    # acting forward equals the oracle on the same weights
    s_t = (np.random.randint(0, 256, size=(84, 84, 4)).astype(np.float32)) * np.float32(1 / 255.0)
    pi, v = net.run_policy_and_value(None, s_t)
    params = net.pop.dump(net.pop.params)
    pi_r, v_r, _, _ = po.forward_backward_numpy(hp.C1, params, s_t[None], geo, [0], [0.0], [0.0], want_grads=False)
    assert hp.rel_err(pi, pi_r[0]) < 1e-3 and abs(v - v_r[0]) < 1e-3 * max(1.0, abs(v_r[0]))
    assert abs(net.run_value(None, s_t) - v) < 1e-6 and np.allclose(net.run_policy(None, s_t), pi)
    # freeze module (0,1): its variables leave get_vars() and are untouched by process()
    fixed = np.zeros((FLAGS.L, FLAGS.M))
    fixed[0, 1] = 1.0
    net.set_fixed_path(fixed)
    assert len(net.get_vars()) == 64 - 2 and net.get_vars_idx()[2:4] == [0, 0]
    th.set_training_stage(0)
    before = net.pop.params.clone()
    ms_before = net.pop.ms.clone()
    diff = th.process(None, 0, None, None, None, None, None, None, FLAGS, None, None)
    assert diff == 20 and np.isfinite(th.last_loss)
    w01 = net.W_conv[0, 1]
    t = net.pop.tensors[w01.index]
    sl = slice(t["offset"], t["offset"] + t["length"])
    assert torch.equal(net.pop.params[sl], before[sl]) and torch.equal(net.pop.ms[sl], ms_before[sl])
    assert not torch.equal(net.pop.params, before)
    assert applier.get_slot(net.W_fc2, "rms").shape == (256, 18)
    assert th._anneal_learning_rate(MAX_TIME_STEP * 2) == 0.0
    net.pop.close()


def test_lstm_network_and_thread_like_the_reference_driver(native_lib):
    """The reference's default (constants.py:31 USE_LSTM = True): GameACPathNetLSTMNetwork through A3CTrainingThread.
    State carry of run_policy_and_value, roll-back of run_value (game_ac_network.py:448-481), reset_state, the extra
    variables in get_vars (:502-520), and one process() call whose training batch is fed in time order from the state
    the rollout started with (a3c_training_thread.py:113-114,184-203) -- checked against the oracle's LSTM."""
    import dataclasses
    from pathnet_gym_b200 import constants, pathnet
    from pathnet_gym_b200.a3c_training_thread import A3CTrainingThread
    from pathnet_gym_b200.constants import GRAD_NORM_CLIP, MAX_TIME_STEP, RMSP_ALPHA, RMSP_EPSILON
    from pathnet_gym_b200.game_ac_network import GameACPathNetLSTMNetwork
    from pathnet_gym_b200.rmsprop_applier import RMSPropApplier
    assert constants.USE_LSTM is True
    FLAGS = make_flags()                            # no use_lstm field: the constant decides, as in the reference
    np.random.seed(0)
    applier = RMSPropApplier(learning_rate=None, decay=RMSP_ALPHA, momentum=0.0, epsilon=RMSP_EPSILON,
                             clip_norm=GRAD_NORM_CLIP, device="/gpu:0")
    th = A3CTrainingThread(0, "", 0, 7.0016e-4, None, applier, MAX_TIME_STEP, device="/gpu:0", FLAGS=FLAGS, task_index=0)
    net = th.local_network
    assert isinstance(net, GameACPathNetLSTMNetwork)
    assert len(net.get_vars()) == 2 * (2 * 10 + 10) + 4 + 2 == len(th.apply_gradients)
    assert net.W_lstm.get_shape() == (512, 1024) and net.b_lstm.get_shape() == (1024,)
    assert net.get_vars_idx()[-6:] == [1, 1, 1, 1, 1, 1]
    net.b_lstm.assign(np.random.default_rng(2).normal(size=1024) * 0.1)      # (the initial bias is all zero)
    pathnet.seed(7)
    geo = pathnet.get_geopath(FLAGS.L, FLAGS.M, FLAGS.N)
    pathnet.geopath_insert(None, net.geopath_update_placeholders_set[0], net.geopath_update_ops_set[0], geo, FLAGS.L, FLAGS.M)
    cfg = hp.C1_LSTM
    params = net.pop.dump(net.pop.params)
    frames = (np.random.randint(0, 256, size=(3, 84, 84, 4)).astype(np.float32)) * np.float32(1 / 255.0)
    # three acting steps carry the state; the oracle runs the same three frames as one sequence from zeros
    outs = [net.run_policy_and_value(None, f) for f in frames[:2]]
    v_boot = net.run_value(None, frames[2])                     # does not advance the state ...
    outs.append(net.run_policy_and_value(None, frames[2]))      # ... so the same frame gives the same value again
    assert abs(outs[2][1] - v_boot) < 1e-6
    st = {}
    pi_r, v_r, _, _ = po.forward_backward_numpy(cfg, params, frames, geo, [0, 0, 0], [0.0] * 3, [0.0] * 3, want_grads=False,
                                                state_out=st)
    for i in range(3):
        assert hp.rel_err(outs[i][0], pi_r[i]) < 1e-3 and abs(outs[i][1] - v_r[i]) < 1e-3 * max(1.0, abs(v_r[i]))
    got = net.lstm_state_out.cpu().numpy()[0]
    assert hp.rel_err(got[0], st[0]) < 1e-3 and hp.rel_err(got[1], st[1]) < 1e-3
    net.reset_state()
    assert float(net.lstm_state_out.abs().max()) == 0.0
    th.set_training_stage(0)
    before = net.pop.params.clone()
    diff = th.process(None, 0, None, None, None, None, None, None, FLAGS, None, None)
    assert diff == 20 and np.isfinite(th.last_loss)
    t = net.pop.tensors[net.W_lstm.index]
    sl = slice(t["offset"], t["offset"] + t["length"])
    assert not torch.equal(net.pop.params[sl], before[sl])      # the cell's weights train
    assert float(net.lstm_state_out.abs().max()) > 0.0          # and the acting state moved on with the rollout
    net.pop.close()


def test_controller_tournament_freeze_reinit(native_lib):
    """doom_pathnet.py:212-293 on a 12-path population."""
    from pathnet_gym_b200.controller import PopulationController
    from pathnet_gym_b200.population import PathNetPopulation
    cfg = hp.to_engine_cfg(hp.TINY)
    P = 12
    pop = PathNetPopulation(cfg, P, 2)
    pop.init_params(3)
    ctl = PopulationController(pop, P, Bcand=3, seed=99)
    ctl.start_task()
    g0 = ctl.genotypes.cpu().numpy().view(np.uint32).copy()
    assert np.array_equal(g0, np.stack([po.get_geopath_masks(cfg.L, cfg.M, cfg.N, 99, 1, p) for p in range(P)]))
    assert np.all(ctl.tournament_step() == -1)                 # nobody has reported a score yet
    scores = torch.arange(P, dtype=torch.float32)
    ctl.report_scores(scores)
    w = ctl.tournament_step()
    g_ref, eff_ref, sc_ref, w_ref = po.tournament(g0, scores.numpy(), np.zeros(cfg.L, np.uint32), cfg.L, cfg.M, cfg.N,
                                                  3, 4, 99, 3)
    assert np.array_equal(w, w_ref) and np.array_equal(ctl.genotypes.cpu().numpy().view(np.uint32), g_ref)
    assert np.array_equal(pop.masks, eff_ref) and float(ctl.scores.max()) == -1000.0
    # task switch: freeze the winner's path, re-initialise the rest, keep the slots
    pop.params.mul_(2.0)
    pop.ms.fill_(0.5)
    fixed = ctl.end_task()
    assert np.array_equal(fixed, po.gates_from_masks(hp.TINY, g_ref[ctl.winner_idx]))
    assert np.array_equal(pop.frozen, po.frozen_flags(hp.TINY, fixed))
    for i, t in enumerate(pop.tensors):
        sl = slice(t["offset"], t["offset"] + t["length"])
        if pop.frozen[i]:
            assert torch.equal(pop.params[sl], ctl.vars_backup[sl] * 2.0)
        else:
            assert torch.equal(pop.params[sl], ctl.vars_backup[sl])
    assert float(pop.ms.min()) == 0.5
    ctl.start_task()                                           # task 2: new genotypes OR the fixed path
    assert np.all((pop.masks & g_ref[ctl.winner_idx][None, :]) == g_ref[ctl.winner_idx][None, :])
    pop.close()


def test_population_rollout_matches_host_bookkeeping(native_lib):
    """PopulationRollout (device bookkeeping) against the reference's host loop restated with the oracle functions,
    on a deterministic synthetic environment: same actions (same Philox draws on the same pi), same R / td, padding
    after terminal steps (a3c_training_thread.py:118-180)."""
    from pathnet_gym_b200.population import PathNetPopulation
    from pathnet_gym_b200.trainer import PopulationRollout
    cfg, P, T = hp.TINY, 4, 6
    pop = PathNetPopulation(hp.to_engine_cfg(cfg), P, T)
    pop.load(pop.params, po.init_params(cfg, 2))
    masks = np.stack([po.get_geopath_masks(cfg.L, cfg.M, cfg.N, 7, 0, p) for p in range(P)])
    pop.set_genotypes(masks)
    dev = pop.device
    gen = torch.Generator(device="cpu").manual_seed(3)
    states = torch.randint(0, 256, (T + 1, P, cfg.H, cfg.W, cfg.C), dtype=torch.uint8, generator=gen).to(dev)
    rewards = (torch.randn(T, P, generator=gen) * 2).to(dev)
    term_at = [T + 5, 2, 4, 0]                                   # path p's episode ends with step term_at[p]
    clock = {"i": 0}

    def env_step(a):
        i = clock["i"]
        clock["i"] += 1
        term = torch.tensor([i >= term_at[q] for q in range(P)], device=dev)
        return states[i + 1], rewards[i], term

    ro = PopulationRollout(pop, T, seed=5)
    frames, actions, td, R, lengths, last = ro.collect(states[0], env_step)
    lengths = lengths.cpu().numpy()
    assert list(lengths) == [min(T, t + 1) for t in term_at]
    acts = actions.cpu().numpy()
    for i in range(T):
        pi, _ = pop.forward(states[i].unsqueeze(1).contiguous())
        want = po.sample_actions(pi.reshape(P, -1).cpu().numpy(), 5, i, id0=0)
        for q in range(P):
            assert acts[q, i] == (want[q] if i < lengths[q] else -1)
    _, vT = pop.forward(last.unsqueeze(1).contiguous())
    boot = np.asarray([float(vT[q, 0]) if term_at[q] >= T else 0.0 for q in range(P)], np.float32)
    Rw, tdw = po.nstep_returns_batch(rewards.t().cpu().numpy(), ro.values.cpu().numpy(), boot, lengths)
    assert np.array_equal(R.cpu().numpy(), Rw) and np.array_equal(td.cpu().numpy(), tdw)
    # and the assembled batch trains: padded transitions carry no loss
    loss = pop.train_step(frames, actions, td, R, 0.01)
    assert torch.isfinite(loss).all()
    pop.close()


def test_population_checkpoint_roundtrip(native_lib, tmp_path):
    from pathnet_gym_b200 import checkpoint as ck
    from pathnet_gym_b200.population import PathNetPopulation
    cfg, P, B = hp.TINY, 2, 3
    pop = PathNetPopulation(hp.to_engine_cfg(cfg), P, B)
    pop.load(pop.params, po.init_params(cfg, 8))
    frames, actions, td, r, masks = hp.synth_inputs(cfg, P, B, 9)
    pop.set_genotypes(masks)
    d = lambda x: torch.from_numpy(np.ascontiguousarray(x)).cuda()
    pop.train_step(d(frames), d(actions), d(td), d(r), 0.01)
    pop.apply_gradients(1e-3)
    path = str(tmp_path / "pop.npz")
    ck.save_population(path, pop, global_t=7)
    want_p, want_s = pop.dump(pop.params), pop.dump(pop.ms)
    pop2 = PathNetPopulation(hp.to_engine_cfg(cfg), P, B)
    got = ck.load_population(path, pop2)
    assert got["global_t"] == 7 and np.array_equal(pop2.masks, masks)
    assert all(np.array_equal(a, b) for a, b in zip(pop2.dump(pop2.params), want_p))
    assert all(np.array_equal(a, b) for a, b in zip(pop2.dump(pop2.ms), want_s))
    # the restored population continues identically
    l1 = pop.train_step(d(frames), d(actions), d(td), d(r), 0.01).cpu().numpy()
    l2 = pop2.train_step(d(frames), d(actions), d(td), d(r), 0.01).cpu().numpy()
    assert np.array_equal(l1, l2)
    pop.close(); pop2.close()
