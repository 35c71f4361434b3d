"""GPU worker for the parameter-server strategy's NVLink transport (launched by tfde_b200.launch: N workers + ps
tasks, one process each). Modes:
  const  every worker pushes K constant gradients; the chief checks the owner shards against the closed form
         w0 - lr * K * sum_r g_r (order-independent), i.e. no update was lost or duplicated by the atomics.
  job    the reference's Estimator flow (mnist_keras_distributed.py:240-283) with ParameterServerStrategy for
         training and MirroredStrategy() for evaluation."""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import numpy as np
import torch

import tfde_b200 as tf
from tfde_b200 import estimator as E


def const(out_dir, K=25):
    st = tf.distribute.experimental.ParameterServerStrategy()
    if st.resolver.is_ps:
        st.run_server()
        return
    with st.scope():
        m = tf.models.build_and_compile_model_b(0.05, seed=3)
    m._bind_optimizer()
    w0 = {v.name: v.numpy().copy() for v in m.variables}
    st.on_train_begin(m)
    r = st.replica_id
    n_workers = len(st.resolver.cluster.training_tasks())
    for k in range(K):
        m.arena.g.fill_((r + 1) * 1e-3)
        st.push(m)
    step = st.global_step()
    st.on_train_end(m)
    out = {"rank": r, "transport": st.transport, "step_seen": step, "device": str(m.device)}
    if st.is_chief:
        total = 0.05 * K * sum((q + 1) * 1e-3 for q in range(n_workers))
        err = 0.0
        for v in m.variables:
            if v.trainable:
                err = max(err, float(np.abs(v.numpy() - (w0[v.name] - total)).max()))
        out.update(max_err=err, expect_delta=total, final_step=st.global_step() if False else None)
    json.dump(out, open(os.path.join(out_dir, f"const_{st.resolver.task.type}{st.resolver.task.index}.json"), "w"))


def job(out_dir):
    tf.cluster.synthesize_tf_config_from_cluster_spec()
    (xtr, ytr), (xte, yte) = tf.data.mnist.load_data(n_train=1024, n_test=256)
    xtr, xte = xtr.reshape(-1, 784) / 255.0, xte.reshape(-1, 784) / 255.0
    ytr, yte = ytr.astype("int").reshape(-1, 1), yte.astype("int").reshape(-1, 1)
    strategy = tf.distribute.experimental.ParameterServerStrategy()
    cfg = E.RunConfig(experimental_distribute=tf.distribute.DistributeConfig(
        train_distribute=strategy, eval_distribute=tf.distribute.MirroredStrategy()),
        session_config=E.session_config_from_env(), model_dir=os.path.join(out_dir, "model"),
        save_summary_steps=10, log_step_count_steps=10, save_checkpoints_steps=500)
    est = E.model_to_estimator(keras_model_builder=lambda: tf.models.build_and_compile_model_b(0.05, seed=3),
                               model_dir=cfg.model_dir, config=cfg)

    def input_fn(x, y, mode):
        ds = tf.data.Dataset.from_tensor_slices((x.astype(np.float32), y))
        if mode == "train":
            return ds.shuffle(1000).repeat().batch(64).prefetch(10)
        return ds.batch(64)

    train_spec = E.TrainSpec(lambda: input_fn(xtr, ytr, "train"), max_steps=60)
    eval_spec = E.EvalSpec(lambda: input_fn(xte, yte, "eval"), steps=None, name="mnist-eval",
                           exporters=[E.FinalExporter("exporter", None)], start_delay_secs=0, throttle_secs=0)
    res = E.train_and_evaluate(est, train_spec, eval_spec)
    role = strategy.resolver.task
    json.dump({"role": role.type, "index": role.index, "global_step": est.global_step,
               "transport": getattr(strategy, "transport", None), "eval": res[0] if res else None},
              open(os.path.join(out_dir, f"job_{role.type}{role.index}.json"), "w"))


if __name__ == "__main__":
    {"const": const, "job": job}[sys.argv[1]](sys.argv[2])
