"""Float-path golden vectors from the REAL reference graph, for the day a TensorFlow with compat.v1 is present.

The reference's arithmetic lives in TensorFlow 1.x (game_ac_network.py, rmsprop_applier.py), which this image does not
have, so the committed float fixtures (float_golden_*.npz) are the oracle's own fp64 outputs and the float path is
"parity unpinned".  This script closes that gap without new work once TensorFlow can be imported (for instance from a
driver-side install under baseline/_ref): it builds the reference's own GameACPathNetNetwork, prepare_loss, tf.gradients
and RMSPropApplier.apply_gradients from the reference sources, feeds the same seeded inputs make_golden.py uses, and
rewrites the fixtures with TensorFlow's outputs (fp32) next to the oracle's fp64 ones.

    python tests/golden/make_tf_golden.py [/root/reference]        # exits 2 with a message when TensorFlow is absent

bench.py --impl reference calls time_reference_graph() through the same probe.
"""
import os
import sys
import time
import types

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def import_tf():
    import importlib
    tf = importlib.import_module("tensorflow.compat.v1")
    tf.disable_v2_behavior()
    return tf


def _flags(cfg, workers=1):
    """The FLAGS fields game_ac_network.py reads (:124-131,158)."""
    return types.SimpleNamespace(task_index=0, L=cfg.L, M=cfg.M, N=cfg.N, worker_hosts_num=workers,
                                 kernel_num=",".join(str(k) for k in cfg.kernels[:cfg.L - 1]),
                                 stride_size=",".join(str(s) for s in cfg.strides[:cfg.L - 1]), use_lstm=False)


def build_reference(ref_dir, cfg, tf):
    """The reference's own classes on its own code path.  The fork hard-codes a 160x120 input and last_lin_num = 1408
    (game_ac_network.py:135,180); for the 84x84 configurations those two literals are the only thing patched, in memory."""
    sys.path.insert(0, ref_dir)
    sys.modules.setdefault("tensorflow", tf)
    src = open(os.path.join(ref_dir, "game_ac_network.py")).read()
    src = src.replace("last_lin_num = 1408", "last_lin_num = %d" % cfg.last_lin)
    src = src.replace("[None, 160, 120, 4]", "[None, %d, %d, %d]" % (cfg.H, cfg.W, cfg.C))
    mod = types.ModuleType("game_ac_network")
    exec(compile(src, os.path.join(ref_dir, "game_ac_network.py"), "exec"), mod.__dict__)
    import constants
    constants.ACTION_SIZEZ = [cfg.A, cfg.A]
    import rmsprop_applier
    return mod, rmsprop_applier


def reference_step(ref_dir, cfg, params, frames_f, gates, actions, td, r, lr=7.0016e-4):
    """pi, v, total_loss, gradients and the updated variables / rms slots of ONE path from the reference graph."""
    from oracle import pathnet_oracle as po
    tf = import_tf()
    tf.reset_default_graph()
    net_mod, rms_mod = build_reference(ref_dir, cfg, tf)
    lr_in = tf.placeholder("float")
    applier = rms_mod.RMSPropApplier(learning_rate=lr_in, decay=po.RMSP_ALPHA, momentum=0.0, epsilon=po.RMSP_EPSILON,
                                     clip_norm=po.GRAD_NORM_CLIP, device="/cpu:0")
    net = net_mod.GameACPathNetNetwork(0, 0, "/cpu:0", _flags(cfg))
    net.prepare_loss(po.ENTROPY_BETA)
    var_refs = [v._ref() for v in net.get_vars()]
    grads = tf.gradients(net.total_loss, var_refs, gate_gradients=False, aggregation_method=None,
                         colocate_gradients_with_ops=False)
    apply_ops = applier.apply_gradients(net.get_vars(), grads)
    a1h = np.zeros((len(actions), cfg.A), np.float32)
    a1h[np.arange(len(actions)), actions] = 1.0
    with tf.Session() as sess:
        sess.run(tf.global_variables_initializer())
        for v, p in zip(net.get_vars(), params):
            sess.run(v.assign(p))
        for l in range(cfg.L):
            for j in range(cfg.M):
                sess.run(net.geopath_set[0][l, j].assign(float(gates[l, j])))
        feed = {net.s: frames_f, net.a: a1h, net.td: td, net.r: r, lr_in: lr}
        pi, v, loss, g = sess.run([net.pi, net.v, net.total_loss, grads], feed_dict=feed)
        sess.run(apply_ops, feed_dict=feed)
        new_vars = sess.run(net.get_vars())
        new_ms = sess.run([applier.get_slot(x, "rms") for x in net.get_vars()])
    return pi, v, loss, g, new_vars, new_ms


def time_reference_graph(ref_dir, cfg, B, n_steps, warmup):
    """(env-steps/s, seconds, threads) of n_steps single-path updates on the reference graph (bench.py --impl reference)."""
    from oracle import pathnet_oracle as po
    tf = import_tf()
    tf.reset_default_graph()
    net_mod, rms_mod = build_reference(ref_dir, cfg, tf)
    lr_in = tf.placeholder("float")
    applier = rms_mod.RMSPropApplier(learning_rate=lr_in, decay=po.RMSP_ALPHA, momentum=0.0, epsilon=po.RMSP_EPSILON,
                                     clip_norm=po.GRAD_NORM_CLIP, device="/cpu:0")
    net = net_mod.GameACPathNetNetwork(0, 0, "/cpu:0", _flags(cfg))
    net.prepare_loss(po.ENTROPY_BETA)
    grads = tf.gradients(net.total_loss, [v._ref() for v in net.get_vars()], gate_gradients=False)
    ops = applier.apply_gradients(net.get_vars(), grads)
    rng = np.random.Generator(np.random.PCG64(1234))
    frames = (rng.integers(0, 256, size=(B, cfg.H, cfg.W, cfg.C)) / 255.0).astype(np.float32)
    a1h = np.eye(cfg.A, dtype=np.float32)[rng.integers(0, cfg.A, size=B)]
    td, r = rng.normal(size=B).astype(np.float32), rng.normal(size=B).astype(np.float32)
    feed = {net.s: frames, net.a: a1h, net.td: td, net.r: r, lr_in: 7.0016e-4}
    with tf.Session() as sess:
        sess.run(tf.global_variables_initializer())
        for _ in range(warmup):
            sess.run(ops, feed_dict=feed)
        t0 = time.perf_counter()
        for _ in range(n_steps):
            sess.run(ops, feed_dict=feed)
        dt = time.perf_counter() - t0
    return B * n_steps / dt, dt, os.cpu_count() or 1


def main():
    ref_dir = sys.argv[1] if len(sys.argv) > 1 else "/root/reference"
    try:
        import_tf()
    except Exception as e:  # noqa: BLE001
        print("TensorFlow with compat.v1 is not importable (%r): the float fixtures stay oracle-generated "
              "(parity unpinned)." % (e,))
        return 2
    from oracle import pathnet_oracle as po
    import make_golden
    for name in ("float_golden_tiny.npz", "float_golden_tiny4.npz", "float_golden_c1.npz"):
        path = os.path.join(HERE, name)
        if not os.path.exists(path):
            continue
        z = dict(np.load(path, allow_pickle=False))
        from tests import helpers as hp
        cfg = hp.cfg_from_golden(z["cfg"])
        frames = po.frames_to_float(z["frames"][0])
        gates = po.gates_from_masks(cfg, z["masks"][0])
        n = len(po.tensor_table(cfg))
        params = [z["param_%d" % i] for i in range(n)]
        pi, v, loss, g, nv, nms = reference_step(ref_dir, cfg, params, frames, gates, z["actions"][0], z["td"][0], z["r"][0])
        z.update({"tf_pi": pi, "tf_v": v, "tf_loss": np.float32(loss)})
        for i in range(n):
            z["tf_grad_%d" % i], z["tf_new_var_%d" % i], z["tf_new_ms_%d" % i] = g[i], nv[i], nms[i]
        np.savez_compressed(path, **z)
        print("pinned", name, "with TensorFlow outputs")
    return 0


if __name__ == "__main__":
    sys.exit(main())
