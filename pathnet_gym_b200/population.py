"""Population-batched PathNet engine: thin Python owner of the device slabs that
calls libpathnet_b200.so through ctypes.  PyTorch is plumbing only (device
memory, streams, torch.distributed for the gradient all-reduce).

The reference evaluates one path per OS process against parameter servers
(doom_pathnet.py:59-77,177-201); here the population is a tensor dimension and
each GPU holds a slice of it.
"""
from __future__ import annotations

import ctypes as C
import math
from dataclasses import dataclass
from typing import List, Optional, Sequence, Tuple

import numpy as np
import torch

from . import _native as nat


@dataclass(frozen=True)
class PathNetConfig:
    """Shape of the super-net; defaults = BASELINE configs C1/C2 (L=3, M=10, N=3, 84x84x4)."""
    L: int = 3
    M: int = 10
    N: int = 3
    H: int = 84
    W: int = 84
    C: int = 4
    kernels: Tuple[int, ...] = (8, 4, 3)       # FLAGS.kernel_num  (doom_pathnet.py:356-357)
    strides: Tuple[int, ...] = (4, 2, 1)       # FLAGS.stride_size (doom_pathnet.py:358-359)
    F: int = 8                                  # feature_num (game_ac_network.py:132)
    hidden: int = 256                           # game_ac_network.py:146
    A: int = 18                                 # max(ACTION_SIZEZ) (game_ac_network.py:152)
    lstm: bool = False                          # GameACPathNetLSTMNetwork (game_ac_network.py:303-521): no /M, BasicLSTMCell(256)


def _ptr(t: Optional[torch.Tensor]):
    return C.c_void_p(t.data_ptr()) if t is not None else C.c_void_p(0)


class PathNetPopulation:
    """P paths x (up to) B_max transitions on one GPU.

    Owns four fp32 slabs laid out in get_vars() order (game_ac_network.py:239-253):
    params, rms (init 1.0, rmsprop_applier.py:34-37), optional momentum (init 0), grads.
    """

    def __init__(self, cfg: PathNetConfig, P: int, B_max: int, device: int = 0, momentum_slot: bool = False,
                 flags: int = 0):
        if not torch.cuda.is_available():
            raise RuntimeError("pathnet_gym_b200 needs a CUDA device (there is no CPU fallback)")
        self.lib = nat.load()
        self.cfg, self.P, self.B_max, self.device_index = cfg, int(P), int(B_max), int(device)
        self.device = torch.device("cuda", self.device_index)
        pc = nat.PnConfig()
        pc.L, pc.M, pc.H, pc.W, pc.C, pc.F, pc.hidden, pc.A = cfg.L, cfg.M, cfg.H, cfg.W, cfg.C, cfg.F, cfg.hidden, cfg.A
        for i in range(cfg.L - 1):
            pc.kernel[i], pc.stride[i] = cfg.kernels[i], cfg.strides[i]
        pc.P, pc.B_max, pc.device, pc.flags = self.P, self.B_max, self.device_index, int(flags) | (nat.PN_FLAG_LSTM if cfg.lstm else 0)
        h = C.c_void_p()
        nat.check(self.lib.pn_create(C.byref(pc), C.byref(h)))
        self._h = h
        n = self.lib.pn_num_tensors(h)
        arr = (nat.PnTensorDesc * n)()
        nat.check(self.lib.pn_tensor_table(h, arr, n), h)
        self.tensors = [dict(offset=int(d.offset), length=int(d.length), layer=int(d.layer), module=int(d.module),
                             kind=int(d.kind)) for d in arr]
        self.shapes = self._shapes()
        self.slab_floats = int(self.lib.pn_slab_floats(h))
        self.last_lin = int(self.lib.pn_last_lin(h))
        with torch.cuda.device(self.device):
            self.params = torch.zeros(self.slab_floats, dtype=torch.float32, device=self.device)
            self.ms = torch.ones(self.slab_floats, dtype=torch.float32, device=self.device)
            self.grads = torch.zeros(self.slab_floats, dtype=torch.float32, device=self.device)
            self.mom = torch.zeros(self.slab_floats, dtype=torch.float32, device=self.device) if momentum_slot else None
        nat.check(self.lib.pn_bind_slabs(h, _ptr(self.params), _ptr(self.ms), _ptr(self.mom), _ptr(self.grads)), h)
        self.masks = None
        self.frozen = np.zeros(n, dtype=np.uint8)

    # -- lifecycle -----------------------------------------------------------
    def close(self):
        if getattr(self, "_h", None):
            torch.cuda.synchronize(self.device)
            self.lib.pn_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def _stream(self):
        return C.c_void_p(torch.cuda.current_stream(self.device).cuda_stream)

    # -- slab layout ---------------------------------------------------------
    def _shapes(self):
        cfg = self.cfg
        shp = []
        for t in self.tensors:
            k = t["kind"]
            if k == 0:
                cin = cfg.C if t["layer"] == 0 else cfg.F
                ks = cfg.kernels[t["layer"]]
                shp.append((ks, ks, cin, cfg.F))
            elif k == 1:
                shp.append((cfg.F,))
            elif k == 2:
                shp.append((t["length"] // cfg.hidden, cfg.hidden))
            elif k == 3:
                shp.append((cfg.hidden,))
            elif k == 4:
                shp.append((cfg.hidden, cfg.A))
            elif k == 5:
                shp.append((cfg.A,))
            elif k == 6:
                shp.append((cfg.hidden, 1))
            elif k == 8:
                shp.append((2 * cfg.hidden, 4 * cfg.hidden))       # W_lstm (basic_lstm_cell/kernel)
            elif k == 9:
                shp.append((4 * cfg.hidden,))                      # b_lstm
            else:
                shp.append((1,))
        return shp

    def view(self, slab: torch.Tensor, i: int) -> torch.Tensor:
        t = self.tensors[i]
        return slab[t["offset"]:t["offset"] + t["length"]].view(self.shapes[i])

    def load(self, slab: torch.Tensor, arrays: Sequence[np.ndarray]):
        assert len(arrays) == len(self.tensors)
        host = slab.detach().cpu().clone()      # keeps the alignment gaps as they are
        for i, a in enumerate(arrays):
            t = self.tensors[i]
            a = np.ascontiguousarray(a, dtype=np.float32)
            assert a.size == t["length"], (i, a.shape, t)
            host[t["offset"]:t["offset"] + t["length"]] = torch.from_numpy(a.reshape(-1))
        slab.copy_(host)

    def dump(self, slab: torch.Tensor) -> List[np.ndarray]:
        host = slab.detach().cpu().numpy()
        return [host[t["offset"]:t["offset"] + t["length"]].reshape(self.shapes[i]).copy()
                for i, t in enumerate(self.tensors)]

    def init_params(self, seed: int = 1):
        """U(-d, d), d = 1/sqrt(fan_in), bias with the weight's d (game_ac_network.py:89-107)."""
        rng = np.random.Generator(np.random.PCG64(seed))
        out, d = [], 1.0
        for t, shp in zip(self.tensors, self.shapes):
            if t["kind"] == 0:
                d = 1.0 / math.sqrt(shp[2] * shp[0] * shp[1])
            elif t["kind"] in (2, 4, 6):
                d = 1.0 / math.sqrt(shp[0])
            elif t["kind"] == 8:                # BasicLSTMCell kernel: glorot_uniform (the scope's default initializer)
                d = math.sqrt(6.0 / (shp[0] + shp[1]))
            if t["kind"] == 9:                  # BasicLSTMCell bias: zeros
                out.append(np.zeros(shp, dtype=np.float32))
                continue
            out.append(rng.uniform(-d, d, size=shp).astype(np.float32))
        self.load(self.params, out)
        return out

    # -- genotype / freeze tables -------------------------------------------
    def set_genotypes(self, masks):
        """masks [P, L] uint32 EFFECTIVE masks (genotype | fixed path)."""
        m = np.ascontiguousarray(np.asarray(masks, dtype=np.uint32).reshape(self.P, self.cfg.L))
        nat.check(self.lib.pn_set_genotypes(self._h, m.ctypes.data_as(C.c_void_p), self._stream()), self._h)
        self.masks = m

    def set_frozen(self, flags):
        f = np.ascontiguousarray(np.asarray(flags, dtype=np.uint8).reshape(len(self.tensors)))
        nat.check(self.lib.pn_set_frozen(self._h, f.ctypes.data_as(C.c_void_p)), self._h)
        self.frozen = f

    def set_fixed_path(self, fixed_path):
        """fixed_path [L, M] of 0/1 -> per-tensor frozen flags (game_ac_network.py:236-271)."""
        fp = np.asarray(fixed_path)
        fl = np.zeros(len(self.tensors), dtype=np.uint8)
        for i, t in enumerate(self.tensors):
            if t["module"] >= 0:
                row = t["layer"] if t["layer"] < self.cfg.L - 1 else -1
                fl[i] = 1 if fp[row, t["module"]] == 1.0 else 0
        self.set_frozen(fl)

    def bind_lstm_state(self, state_in: Optional[torch.Tensor], state_out: Optional[torch.Tensor] = None):
        """LSTM variant: every path's cell state (c, h) at the start of its sequence, [P, 2, 256] fp32 on the device
        (None = zeros, reset_state), and where the state after the last step is written (None = nowhere).  The
        tensors are borrowed until the next call."""
        for t in (state_in, state_out):
            if t is not None and (t.dtype != torch.float32 or tuple(t.shape) != (self.P, 2, self.cfg.hidden)
                                  or not t.is_contiguous() or t.device != self.device):
                raise ValueError("LSTM state must be a contiguous fp32 [P=%d, 2, %d] tensor on %s" % (self.P, self.cfg.hidden, self.device))
        self._lstm_state = (state_in, state_out)        # keep them alive while the library holds the pointers
        nat.check(self.lib.pn_lstm_bind_state(self._h, _ptr(state_in), _ptr(state_out)), self._h)

    # -- hot path -------------------------------------------------------------
    def _check_frames(self, frames: torch.Tensor):
        cfg = self.cfg
        if frames.dtype not in (torch.uint8, torch.float32):
            raise TypeError("frames must be uint8 (k means k/255) or float32")
        if frames.dim() != 5 or frames.shape[0] != self.P or tuple(frames.shape[2:]) != (cfg.H, cfg.W, cfg.C):
            raise ValueError("frames must be [P=%d, B, %d, %d, %d], got %s" % (self.P, cfg.H, cfg.W, cfg.C, tuple(frames.shape)))
        if not frames.is_contiguous():
            raise ValueError("frames must be contiguous")
        return int(frames.shape[1]), 1 if frames.dtype == torch.uint8 else 0

    def forward(self, frames: torch.Tensor):
        """pi [P,B,A], v [P,B] for device frames (run_policy_and_value for all paths)."""
        B, u8 = self._check_frames(frames)
        pi = torch.empty(self.P, B, self.cfg.A, dtype=torch.float32, device=self.device)
        v = torch.empty(self.P, B, dtype=torch.float32, device=self.device)
        nat.check(self.lib.pn_forward(self._h, _ptr(frames), u8, B, _ptr(pi), _ptr(v), self._stream()), self._h)
        return pi, v

    def train_step(self, frames, actions, td, r, entropy_beta=0.01, want_outputs=False):
        """forward + A3C loss + backward for every path; grads slab <- sum over local paths."""
        B, u8 = self._check_frames(frames)
        for name, t, dt in (("actions", actions, torch.int32), ("td", td, torch.float32), ("r", r, torch.float32)):
            if t.dtype != dt or tuple(t.shape) != (self.P, B) or not t.is_contiguous() or t.device != frames.device:
                raise ValueError("%s must be a contiguous %s [P, B] tensor on the frames' device" % (name, dt))
        loss = torch.empty(self.P, dtype=torch.float32, device=self.device)
        pi = v = None
        if want_outputs:
            pi = torch.empty(self.P, B, self.cfg.A, dtype=torch.float32, device=self.device)
            v = torch.empty(self.P, B, dtype=torch.float32, device=self.device)
        nat.check(self.lib.pn_train_step(self._h, _ptr(frames), u8, _ptr(actions), _ptr(td), _ptr(r), B,
                                         float(entropy_beta), _ptr(loss), _ptr(pi), _ptr(v), self._stream()), self._h)
        return (loss, pi, v) if want_outputs else loss

    def train_step_host(self, frames, actions, td, r, entropy_beta=0.01, loss_out: Optional[torch.Tensor] = None):
        """Same, for HOST tensors (pinned is fastest); copies happen inside the C call."""
        B, u8 = self._check_frames(frames)
        if loss_out is None:
            loss_out = torch.empty(self.P, dtype=torch.float32)
        nat.check(self.lib.pn_train_step_host(self._h, _ptr(frames), u8, _ptr(actions), _ptr(td), _ptr(r), B,
                                              float(entropy_beta), _ptr(loss_out), self._stream()), self._h)
        return loss_out

    def stage_host(self, slot, frames, actions, td, r, copy_stream: Optional[torch.cuda.Stream] = None):
        """Enqueue the H2D copies of one batch (pinned host tensors) into staging slot 0/1 on `copy_stream`."""
        B, u8 = self._check_frames(frames)
        cs = copy_stream if copy_stream is not None else torch.cuda.current_stream(self.device)
        nat.check(self.lib.pn_stage_inputs_host(self._h, int(slot), _ptr(frames), u8, _ptr(actions), _ptr(td), _ptr(r), B,
                                                C.c_void_p(cs.cuda_stream)), self._h)

    def stage_rollout_host(self, slot, planes, actions, td, r, copy_stream: Optional[torch.cuda.Stream] = None):
        """Like stage_host for one rollout per path: `planes` [P, B + C - 1, H, W] u8 (pinned host) holds the distinct
        planes of the B frame stacks (game_state.py:62-66,75-81); the stacks are rebuilt on the device."""
        cfg = self.cfg
        if planes.dtype != torch.uint8 or planes.dim() != 4 or planes.shape[0] != self.P or \
                tuple(planes.shape[2:]) != (cfg.H, cfg.W) or not planes.is_contiguous():
            raise ValueError("planes must be a contiguous uint8 [P=%d, B + %d, %d, %d] tensor" % (self.P, cfg.C - 1, cfg.H, cfg.W))
        B = int(planes.shape[1]) - (cfg.C - 1)
        cs = copy_stream if copy_stream is not None else torch.cuda.current_stream(self.device)
        nat.check(self.lib.pn_stage_rollout_host(self._h, int(slot), _ptr(planes), _ptr(actions), _ptr(td), _ptr(r), B,
                                                 C.c_void_p(cs.cuda_stream)), self._h)

    def train_step_staged(self, slot, entropy_beta=0.01, loss_out: Optional[torch.Tensor] = None):
        """Run the step on a staged slot; the D2H copy of the losses is enqueued, not awaited."""
        nat.check(self.lib.pn_train_step_staged(self._h, int(slot), float(entropy_beta), _ptr(loss_out), self._stream()),
                  self._h)
        return loss_out

    def allreduce_grads(self):
        """Sum the shared-module gradient slab over the GPUs that hold population slices."""
        import torch.distributed as dist
        if dist.is_available() and dist.is_initialized() and dist.get_world_size() > 1:
            dist.all_reduce(self.grads, op=dist.ReduceOp.SUM)

    def lin_grads_region(self):
        """(offset, count) in floats of the W_lin/b_lin gradients inside the gradient slab."""
        import ctypes as C
        off, cnt = C.c_int64(0), C.c_int64(0)
        nat.check(self.lib.pn_lin_grads_region(self._h, C.byref(off), C.byref(cnt)), self._h)
        return int(off.value), int(cnt.value)

    def peer_buffer_bytes(self) -> int:
        return int(self.lib.pn_peer_buffer_bytes(self._h))

    def peer_check(self, my_peer_buf_ptr):
        """Raises PathNetError if an exchange kernel of this rank gave up waiting for a peer (synchronises)."""
        nat.check(self.lib.pn_peer_check(self._h, C.c_void_p(int(my_peer_buf_ptr))), self._h)

    def peer_sum_small(self, peer_ptrs, rank, world, step):
        """Conv + head gradient regions <- sum over the node's GPUs through peer memory (pn_peer_sum_small)."""
        import ctypes as C
        arr = (C.c_uint64 * len(peer_ptrs))(*[int(p) for p in peer_ptrs])
        nat.check(self.lib.pn_peer_sum_small(self._h, arr, int(rank), int(world), int(step), self._stream()), self._h)

    def peer_sum_part(self, peer_ptrs, rank, world, step, part):
        """One part of peer_sum_small (pn_peer_sum_part): 1 = the regions final before the first conv layer's weight
        gradient, 2 = that layer's own tensors."""
        import ctypes as C
        arr = (C.c_uint64 * len(peer_ptrs))(*[int(p) for p in peer_ptrs])
        nat.check(self.lib.pn_peer_sum_part(self._h, arr, int(rank), int(world), int(step), int(part), self._stream()), self._h)

    def peer_allreduce_lin(self, peer_slabs, peer_ptrs, rank, world, step):
        """Linear-module gradient region <- sum over the node's GPUs, in place through peer memory (pn_peer_allreduce_lin)."""
        import ctypes as C
        a = (C.c_uint64 * len(peer_slabs))(*[int(p) for p in peer_slabs])
        b = (C.c_uint64 * len(peer_ptrs))(*[int(p) for p in peer_ptrs])
        nat.check(self.lib.pn_peer_allreduce_lin(self._h, a, b, int(rank), int(world), int(step), self._stream()), self._h)

    def rebind_grads(self, grads: torch.Tensor):
        """Use `grads` (fp32, slab_floats, same device; e.g. symmetric memory) as the gradient slab from now on."""
        assert grads.is_cuda and grads.dtype == torch.float32 and grads.numel() == self.slab_floats and grads.is_contiguous()
        nat.check(self.lib.pn_bind_slabs(self._h, _ptr(self.params), _ptr(self.ms), _ptr(self.mom), _ptr(grads)), self._h)
        self.grads = grads

    def stream_wait_lin_grads(self, stream: "torch.cuda.Stream"):
        """`stream` waits until the latest train_step has written the W_lin/b_lin gradients."""
        import ctypes as C
        nat.check(self.lib.pn_stream_wait_lin_grads(self._h, C.c_void_p(stream.cuda_stream)), self._h)

    def stream_wait_early_grads(self, stream: "torch.cuda.Stream"):
        """`stream` waits until the latest train_step has written every gradient except the first conv layer's; returns
        the number of floats at the start of the slab that are NOT covered (that layer's tensors)."""
        import ctypes as C
        late = C.c_int64(0)
        nat.check(self.lib.pn_stream_wait_early_grads(self._h, C.c_void_p(stream.cuda_stream), C.byref(late)), self._h)
        return int(late.value)

    def apply_gradients(self, lr, decay=0.99, momentum=0.0, epsilon=0.1, clip_norm=40.0):
        """Per-tensor clip_by_norm + RMSProp on the non-frozen tensors (rmsprop_applier.py:79-106)."""
        nat.check(self.lib.pn_rmsprop_apply(self._h, float(lr), float(decay), float(momentum), float(epsilon),
                                            float(clip_norm), self._stream()), self._h)

    def apply_gradients_range(self, first_tensor, n_tensors, lr, decay=0.99, momentum=0.0, epsilon=0.1, clip_norm=40.0,
                              complement=False):
        """apply_gradients for a contiguous range of the tensor table (or for everything outside it)."""
        nat.check(self.lib.pn_rmsprop_apply_range(self._h, int(first_tensor), int(n_tensors), 1 if complement else 0,
                                                  float(lr), float(decay), float(momentum), float(epsilon),
                                                  float(clip_norm), self._stream()), self._h)

    def lin_tensor_range(self):
        """(first, count) of the W_lin / b_lin rows of the tensor table."""
        off, cnt = self.lin_grads_region()
        idx = [i for i, t in enumerate(self.tensors) if off <= t["offset"] < off + cnt]
        return idx[0], len(idx)

    def reinit_unfrozen(self, backup: torch.Tensor):
        """Every non-frozen tensor <- backup (doom_pathnet.py:289-293)."""
        assert backup.dtype == torch.float32 and backup.numel() == self.slab_floats and backup.is_cuda
        nat.check(self.lib.pn_reinit_unfrozen(self._h, _ptr(backup), self._stream()), self._h)

    def relu_decisions(self, layer: int, B: int) -> np.ndarray:
        """Test aid: bool array of the last step's ReLU decisions in `layer`: [P, B, M, OH*OW, F] for a conv layer,
        [P, B, M, hidden] for the linear layer (rows of inactive modules are undefined)."""
        cfg = self.cfg
        lin = layer == cfg.L - 1
        if lin:
            per = cfg.hidden
        else:
            ih, iw = cfg.H, cfg.W
            for l in range(layer + 1):
                ih, iw = (ih - cfg.kernels[l]) // cfg.strides[l] + 1, (iw - cfg.kernels[l]) // cfg.strides[l] + 1
            per = ih * iw
        buf = np.empty(self.P * B * cfg.M * per, dtype=np.uint8)
        n = int(self.lib.pn_debug_read_relu_masks(self._h, int(layer), int(B), buf.ctypes.data_as(C.c_void_p), buf.size))
        if n < 0:
            nat.check(n, self._h)
        buf = buf.reshape(self.P, B, cfg.M, per)
        if lin:
            return buf != 0
        return ((buf[..., None] >> np.arange(cfg.F, dtype=np.uint8)) & 1).astype(bool)

    # -- measurement support ---------------------------------------------------
    def engine_paths(self) -> str:
        """Which engine (tcgen05 or CUDA cores) every layer of this handle runs on."""
        buf = C.create_string_buffer(512)
        self.lib.pn_engine_paths(self._h, buf, 512)
        return buf.value.decode()

    @property
    def launch_count(self) -> int:
        return int(self.lib.pn_launch_count(self._h))

    def profile_enable(self, on: bool = True):
        nat.check(self.lib.pn_profile_enable(self._h, 1 if on else 0), self._h)

    def profile_read(self):
        """{kernel name: (avg ms per launch, launches)} from CUDA events on the launch stream."""
        arr = (nat.PnProfileEntry * 64)()
        n = nat.check(self.lib.pn_profile_read(self._h, arr, 64), self._h)
        return {arr[i].name.decode(): (arr[i].total_ms / max(arr[i].launches, 1), int(arr[i].launches)) for i in range(n)}

    @property
    def workspace_bytes(self) -> int:
        return int(self.lib.pn_workspace_bytes(self._h))


# ---------------------------------------------------------------------------
# genotype ops on device bitmasks
# ---------------------------------------------------------------------------
def geopath_init(n_paths, L, M, N, seed, epoch, path_id0=0, device=0) -> torch.Tensor:
    lib = nat.load()
    dev = torch.device("cuda", device)
    masks = torch.zeros(n_paths, L, dtype=torch.int32, device=dev)
    st = C.c_void_p(torch.cuda.current_stream(dev).cuda_stream)
    nat.check(lib.pn_geopath_init(_ptr(masks), n_paths, path_id0, L, M, N, int(seed), int(epoch), st))
    return masks


def sample_actions(pi: torch.Tensor, seed, epoch, id0=0) -> torch.Tensor:
    """choose_action (a3c_training_thread.py:89-90) for every row of pi [..., A]: int32 actions, one Philox draw per row."""
    lib = nat.load()
    assert pi.is_cuda and pi.dtype == torch.float32 and pi.is_contiguous()
    A = pi.shape[-1]
    n = pi.numel() // A
    out = torch.empty(pi.shape[:-1], dtype=torch.int32, device=pi.device)
    st = C.c_void_p(torch.cuda.current_stream(pi.device).cuda_stream)
    nat.check(lib.pn_sample_actions(_ptr(pi), n, A, int(seed), int(epoch), int(id0), _ptr(out), st))
    return out


def nstep_returns(rewards: torch.Tensor, values: torch.Tensor, bootstrap: torch.Tensor, lengths=None, gamma=0.99):
    """Batch assembly of A3CTrainingThread.process (a3c_training_thread.py:136,155-180) for [P, B] rollouts:
    returns (R, td) fp32 [P, B]; entries at or beyond lengths[p] are padding (0)."""
    lib = nat.load()
    P, B = rewards.shape
    for t in (rewards, values, bootstrap):
        assert t.is_cuda and t.dtype == torch.float32 and t.is_contiguous()
    assert values.shape == (P, B) and bootstrap.shape == (P,)
    if lengths is not None:
        assert lengths.is_cuda and lengths.dtype == torch.int32 and lengths.shape == (P,)
    R = torch.empty(P, B, dtype=torch.float32, device=rewards.device)
    td = torch.empty_like(R)
    st = C.c_void_p(torch.cuda.current_stream(rewards.device).cuda_stream)
    nat.check(lib.pn_nstep_returns(_ptr(rewards), _ptr(values), _ptr(bootstrap), _ptr(lengths), P, B, float(gamma),
                                   _ptr(R), _ptr(td), st))
    return R, td


def preprocess_frames_(obs: torch.Tensor, frames: torch.Tensor, reset: Optional[torch.Tensor] = None) -> torch.Tensor:
    """GameState._process_frame + stack update (game_state.py:37-51,75-81) for n environments, in place on `frames`:
    obs [n, Hin, Win, 3] u8 (BGR order as cv2 sees it) -> gray -> bilinear resize -> newest channel of the u8 stack
    frames [n, H, W, C]; reset [n] u8/bool: fill all C channels (GameState.reset)."""
    lib = nat.load()
    assert obs.is_cuda and obs.dtype == torch.uint8 and obs.is_contiguous() and obs.dim() == 4 and obs.shape[3] == 3
    assert frames.is_cuda and frames.dtype == torch.uint8 and frames.is_contiguous() and frames.dim() == 4
    n, Hin, Win, _ = obs.shape
    assert frames.shape[0] == n
    rs = None
    if reset is not None:
        rs = reset.to(torch.uint8).contiguous()
        assert rs.is_cuda and rs.shape == (n,)
    st = C.c_void_p(torch.cuda.current_stream(obs.device).cuda_stream)
    nat.check(lib.pn_preprocess_frames(_ptr(obs), n, Hin, Win, _ptr(frames), frames.shape[1], frames.shape[2],
                                       frames.shape[3], _ptr(rs), st))
    return frames


def mutate_(masks: torch.Tensor, M, N, seed, epoch, path_id0=0) -> torch.Tensor:
    lib = nat.load()
    assert masks.is_cuda and masks.dtype == torch.int32 and masks.is_contiguous()
    st = C.c_void_p(torch.cuda.current_stream(masks.device).cuda_stream)
    nat.check(lib.pn_mutate(_ptr(masks), masks.shape[0], path_id0, masks.shape[1], M, N, int(seed), int(epoch), st))
    return masks


def tournament_(genotypes: torch.Tensor, scores: torch.Tensor, fixed: torch.Tensor, M, N, Bcand, n_groups, seed, epoch):
    """In-place controller pass; returns (effective masks, winners)."""
    lib = nat.load()
    P, L = genotypes.shape
    assert genotypes.dtype == torch.int32 and scores.dtype == torch.float32 and fixed.dtype == torch.int32
    eff = torch.empty_like(genotypes)
    winners = torch.full((max(n_groups, 1),), -1, dtype=torch.int32, device=genotypes.device)
    st = C.c_void_p(torch.cuda.current_stream(genotypes.device).cuda_stream)
    nat.check(lib.pn_tournament(_ptr(genotypes), _ptr(scores), _ptr(fixed), _ptr(eff), _ptr(winners), P, L, M, N,
                                Bcand, n_groups, int(seed), int(epoch), st))
    return eff, winners[:n_groups]


def rmsprop_apply_noalloc(params, ms, grads, tensors, lr, decay=0.99, momentum=0.0, epsilon=0.1, clip_norm=40.0,
                          mom=None, frozen=None):
    """The clipped RMSProp update without a handle and without any device allocation, for hosts that
    cannot call cudaMalloc when they update (fully reserved GPU, no tolerance for cudaMalloc's
    device-wide synchronisation, plugins that may not own device memory).  `tensors` is a list of
    dicts with `offset` and `length` (floats into the slabs), e.g. PathNetPopulation.tensors.
    Limits: 128 tensors, 16384 chunks of 2048 floats; larger inputs raise PathNetError.  One scratch
    array per device is shared by all calls: do not let two calls overlap on the same GPU."""
    lib = nat.load()
    n = len(tensors)
    arr = (nat.PnTensorDesc * max(n, 1))()
    for i, t in enumerate(tensors):
        arr[i].offset, arr[i].length = int(t["offset"]), int(t["length"])
        arr[i].layer, arr[i].module, arr[i].kind = int(t.get("layer", 0)), int(t.get("module", -1)), int(t.get("kind", 0))
    fz = None
    if frozen is not None:
        fz = np.ascontiguousarray(np.asarray(frozen, dtype=np.uint8))
        assert fz.size == n
    with torch.cuda.device(params.device):
        st = C.c_void_p(torch.cuda.current_stream(params.device).cuda_stream)
        nat.check(lib.pn_rmsprop_apply_noalloc(_ptr(params), _ptr(ms), _ptr(mom), _ptr(grads), arr,
                                               fz.ctypes.data_as(C.c_void_p) if fz is not None else C.c_void_p(0), n,
                                               float(lr), float(decay), float(momentum), float(epsilon),
                                               float(clip_norm), st))
