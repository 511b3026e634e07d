# -*- coding: utf-8 -*-
"""Drop-in for a3c_training_thread.py: rollout bookkeeping, n-step returns, learning-rate anneal and
the one training call per update (a3c_training_thread.py:25-245).  The environment adapter
(game_state.py, gym / ALE / ViZDoom) is out of scope: a `game_state` object with the reference's
interface (s_t, reward, terminal, process(action), update(), reset()) is injected; the default is a
synthetic frame source."""
import time

import numpy as np
import torch

from .constants import ACTION_SIZEZ, ENTROPY_BETA, GAMMA, LOCAL_T_MAX, ROMZ, USE_LSTM
from .game_ac_network import GameACPathNetLSTMNetwork, GameACPathNetNetwork

LOG_INTERVAL = 100
PERFORMANCE_LOG_INTERVAL = 1000


class SyntheticGameState(object):
    """Stand-in for GameState (game_state.py:19-84): random k/255 frames, 4-frame stack, random rewards."""

    def __init__(self, rand_seed, H, W, episode_len=50):
        self.rng = np.random.Generator(np.random.PCG64(rand_seed))
        self.H, self.W, self.episode_len = H, W, episode_len
        self.reset()

    def _frame(self):
        return (self.rng.integers(0, 256, size=(self.H, self.W, 1)).astype(np.float32)) * np.float32(1.0 / 255.0)

    def reset(self):
        x = self._frame()
        self.s_t = np.concatenate([x, x, x, x], axis=2)
        self.reward, self.terminal, self.t = 0, False, 0

    def process(self, action):
        self.t += 1
        self.reward = float(self.rng.integers(-2, 3))
        self.terminal = self.t >= self.episode_len
        self.s_t1 = np.append(self.s_t[:, :, 1:], self._frame(), axis=2)

    def update(self):
        self.s_t = self.s_t1

    def close_env(self):
        pass


class A3CTrainingThread(object):
    def __init__(self, thread_index, global_network, training_stage, initial_learning_rate, learning_rate_input,
                 grad_applier, max_global_time_step, device, FLAGS="", task_index=""):
        self.training_stage = training_stage
        self.thread_index = thread_index
        self.task_index = task_index
        self.learning_rate_input = learning_rate_input
        self.max_global_time_step = max_global_time_step
        # the reference overwrites FLAGS.use_lstm with the constant (a3c_training_thread.py:46); a caller that set the
        # field on its FLAGS object before is respected here so that both networks can be driven in one process
        if not hasattr(FLAGS, "use_lstm"):
            FLAGS.use_lstm = USE_LSTM
        self.use_lstm = bool(FLAGS.use_lstm)
        if self.use_lstm:
            self.local_network = GameACPathNetLSTMNetwork(training_stage, thread_index, device, FLAGS)
        else:
            self.local_network = GameACPathNetNetwork(training_stage, thread_index, device, FLAGS)
        self.local_network.prepare_loss(ENTROPY_BETA)
        # tf.gradients + apply ops for the full variable list (fixed_path is all-zero here,
        # a3c_training_thread.py:55-64)
        self.grad_applier = grad_applier
        self.gradients = None
        self.apply_gradients = grad_applier.apply_gradients(self.local_network.get_vars(), self.gradients)
        self.local_t = 0
        self.initial_learning_rate = initial_learning_rate
        self.episode_reward = 0
        self.prev_local_t = 0
        self.game_state = None
        self.game_state_factory = lambda seed, rom, **kw: SyntheticGameState(
            seed, self.local_network.cfg.H, self.local_network.cfg.W)
        self.start_time = time.time()

    def set_training_stage(self, training_stage):
        """a3c_training_thread.py:75-81."""
        self.training_stage = training_stage
        self.local_network.set_training_stage(training_stage)
        if training_stage == 1 and self.game_state is not None:
            self.game_state.close_env()
        self.game_state = self.game_state_factory(113 * int(self.task_index or 0), ROMZ[training_stage], display=False,
                                                  no_op_max=ACTION_SIZEZ[training_stage], task_index=self.task_index)

    def _anneal_learning_rate(self, global_time_step):
        """a3c_training_thread.py:83-87."""
        learning_rate = self.initial_learning_rate * (self.max_global_time_step - global_time_step) / self.max_global_time_step
        if learning_rate < 0.0:
            learning_rate = 0.0
        return learning_rate

    def choose_action(self, pi_values):
        """a3c_training_thread.py:89-90."""
        p = np.asarray(pi_values, dtype=np.float64)
        return np.random.choice(range(len(p)), p=p / p.sum())

    def set_start_time(self, start_time):
        self.start_time = start_time

    def process(self, sess, global_t, summary_writer, summary_op, score_input, score_ph, score_ops, geopath, FLAGS,
                score_set_ph, score_set_ops):
        """One A3C update of this worker's path (a3c_training_thread.py:102-245). `score_ops` /
        `score_set_ops` may be callables taking the episode reward (they replace sess.run(assign))."""
        states, actions, rewards, values = [], [], [], []
        terminal_end = False
        start_local_t = self.local_t
        net = self.local_network
        if self.use_lstm:
            start_lstm_state = net.lstm_state_out            # a3c_training_thread.py:113-114
        for i in range(LOCAL_T_MAX):
            pi_, value_ = net.run_policy_and_value(sess, self.game_state.s_t)
            action = self.choose_action(pi_)
            states.append(self.game_state.s_t)
            actions.append(action)
            values.append(value_)
            self.game_state.process(action)
            reward, terminal = self.game_state.reward, self.game_state.terminal
            self.episode_reward += reward
            rewards.append(np.clip(reward, -1, 1))
            self.local_t += 1
            self.game_state.update()
            if terminal:
                terminal_end = True
                for op in (score_ops, score_set_ops):
                    if callable(op):
                        op(self.episode_reward)
                self.episode_reward = 0
                self.game_state.reset()
                if self.use_lstm:
                    net.reset_state()                        # a3c_training_thread.py:148-149
                break
        R = 0.0
        if not terminal_end:
            R = net.run_value(sess, self.game_state.s_t)
        batch_si, batch_a, batch_td, batch_R = [], [], [], []
        for ai, ri, si, Vi in zip(reversed(actions), reversed(rewards), reversed(states), reversed(values)):
            R = ri + GAMMA * R
            batch_si.append(si)
            batch_a.append(ai)
            batch_td.append(R - Vi)
            batch_R.append(R)
        cur_learning_rate = self._anneal_learning_rate(global_t)
        if self.use_lstm:
            # the cell consumes the rollout in time order from the state it started with (a3c_training_thread.py:184-203)
            batch_si.reverse(); batch_a.reverse(); batch_td.reverse(); batch_R.reverse()
            net.pop.bind_lstm_state(start_lstm_state, None)
        # op selection = freeze mask (a3c_training_thread.py:205-209); then ONE call does
        # forward + loss + backward, and the selected apply ops run
        var_idx = net.get_vars_idx()
        ops = [self.apply_gradients[i] for i in range(len(var_idx)) if var_idx[i] == 1.0]
        B = len(batch_a)
        dev = net.pop.device
        P = net.P
        frames = torch.as_tensor(np.stack(batch_si).astype(np.float32)).to(dev).view(1, B, net.cfg.H, net.cfg.W, net.cfg.C)
        a = torch.as_tensor(np.asarray(batch_a, dtype=np.int32)).to(dev).view(1, B)
        td = torch.as_tensor(np.asarray(batch_td, dtype=np.float32)).to(dev).view(1, B)
        r = torch.as_tensor(np.asarray(batch_R, dtype=np.float32)).to(dev).view(1, B)
        if P != 1:
            raise ValueError("A3CTrainingThread drives one path; use trainer.PopulationTrainer for P > 1")
        self.last_loss = float(net.pop.train_step(frames, a, td, r, net.entropy_beta)[0])
        self.grad_applier.run(ops, cur_learning_rate)
        if (self.task_index == 0) and (self.local_t - self.prev_local_t >= PERFORMANCE_LOG_INTERVAL):
            self.prev_local_t += PERFORMANCE_LOG_INTERVAL
            elapsed_time = time.time() - self.start_time
            steps_per_sec = global_t / max(elapsed_time, 1e-9)
            print("### Performance : {} STEPS in {:.0f} sec. {:.0f} STEPS/sec. {:.2f}M STEPS/hour".format(
                global_t, elapsed_time, steps_per_sec, steps_per_sec * 3600 / 1000000.))
        return self.local_t - start_local_t
