# -*- coding: utf-8 -*-
"""Hyper-parameters of the reference (constants.py:4-32).  The values are part of the numerical
contract of the hot path (RMSProp decay / epsilon, clip norm, gamma, entropy beta, t_max)."""
import time

LOCAL_T_MAX = 20            # rollout length per update                     (constants.py:4)
RMSP_ALPHA = 0.99           # RMSProp decay                                 (constants.py:5)
RMSP_EPSILON = 0.1          # RMSProp epsilon (inside the sqrt)             (constants.py:6)
CHECKPOINT_DIR = "checkpoints"
INITIAL_ALPHA_LOW = 1e-4    # log-uniform learning-rate range               (constants.py:8-9)
INITIAL_ALPHA_HIGH = 1e-2
PARALLEL_SIZE = 8
ROMZ = ["MsPacman-v0", "Alien-v0"]
ACTION_SIZEZ = [9, 18]      # both tasks share one head of width max()      (game_ac_network.py:152)
ACTION_SPACE_TYPE = "full"
GYM_MONITOR_DIR = "./data/gym/experiment-" + str(int(time.time()))
INITIAL_ALPHA_LOG_RATE = 0.4226   # -> lr0 = 7.0016e-4                      (doom_pathnet.py:46-57)
GAMMA = 0.99
ENTROPY_BETA = 0.01
MAX_TIME_STEP = 4 * 10 ** 6
GRAD_NORM_CLIP = 40.0
NUM_GPUS = 3
USE_GPU = True              # the reference ships False (constants.py:30); this build is GPU-only
USE_LSTM = True             # as the reference ships it (constants.py:31): GameACPathNetLSTMNetwork
USE_PATHNET = True
