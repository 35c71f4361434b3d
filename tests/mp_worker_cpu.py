"""CPU/gloo worker for the distributed plumbing tests (BASELINE.json config 1: "Keras MNIST CNN,
MultiWorkerMirroredStrategy-equivalent, 2 workers via TF_CONFIG on localhost CPU/gloo").
Mode is argv[1]; results are written as JSON/NPZ into argv[2] (one file per task)."""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
os.environ.setdefault("TFDE_DEVICE", "cpu")

import numpy as np
import torch

import tfde_b200 as tf
from tfde_b200 import estimator as E


def data(n=512):
    (xtr, ytr), (xte, yte) = tf.data.mnist.load_data(n_train=n, n_test=128)
    return xtr, ytr.astype(np.int64), xte, yte.astype(np.int64)


def mwms(out_dir):
    strategy = tf.distribute.experimental.MultiWorkerMirroredStrategy()
    xtr, ytr, _, _ = data()
    per = 16
    gbs = per * strategy.num_replicas_in_sync
    with strategy.scope():
        ds = tf.data.Dataset.from_tensor_slices((xtr[..., None], ytr)).batch(gbs)   # no shuffle: same order everywhere
        opts = tf.data.Options()
        opts.experimental_distribute.auto_shard_policy = tf.data.experimental.AutoShardPolicy.DATA
        ds = ds.with_options(opts)
        model = tf.models.build_and_compile_cnn_model(learning_rate=0.05, seed=100 + strategy.rank)  # chief's init must win
    hist = model.fit(ds, epochs=2, steps_per_epoch=4, verbose=0)
    if strategy.is_chief:
        model.save_weights(os.path.join(out_dir, "chief_weights.npz"))
    np.savez(os.path.join(out_dir, f"weights_rank{strategy.rank}.npz"), **model.named_weights())
    ev = model.evaluate(xtr[:100, ..., None], ytr[:100], batch_size=50, return_dict=True)
    json.dump({"rank": strategy.rank, "world": strategy.num_replicas_in_sync, "chief": strategy.is_chief,
               "loss": hist.history["loss"], "accuracy": hist.history["accuracy"], "eval": ev},
              open(os.path.join(out_dir, f"result_rank{strategy.rank}.json"), "w"))
    strategy.shutdown()


def ps_job(out_dir):
    """mnist_keras_distributed.py flow: CLUSTER_SPEC shim -> PS train strategy -> train_and_evaluate."""
    tf.cluster.synthesize_tf_config_from_cluster_spec()
    xtr, ytr, xte, yte = data(256)
    xtr, xte = xtr / 255.0, xte / 255.0
    ytr, yte = ytr.reshape(-1, 1), yte.reshape(-1, 1)
    strategy = tf.distribute.experimental.ParameterServerStrategy()
    cfg = E.RunConfig(experimental_distribute=tf.distribute.DistributeConfig(train_distribute=strategy),
                      session_config=E.session_config_from_env(), model_dir=os.path.join(out_dir, "model"),
                      save_summary_steps=10, log_step_count_steps=10, save_checkpoints_steps=500)
    est = E.model_to_estimator(keras_model_builder=lambda: tf.models.build_and_compile_model_b(0.05, seed=3),
                               model_dir=cfg.model_dir, config=cfg)

    def input_fn(x, y, mode):
        ds = tf.data.Dataset.from_tensor_slices((x, y))
        if mode == "train":
            return ds.shuffle(1000).repeat().batch(32).prefetch(10)
        return ds.batch(32)

    train_spec = E.TrainSpec(lambda: input_fn(xtr, ytr, "train"), max_steps=30)
    eval_spec = E.EvalSpec(lambda: input_fn(xte, yte, "eval"), steps=None, name="mnist-eval",
                           exporters=[E.FinalExporter("exporter", None)], start_delay_secs=0, throttle_secs=0)
    res = E.train_and_evaluate(est, train_spec, eval_spec)
    role = strategy.resolver.task
    json.dump({"role": role.type, "index": role.index, "global_step": est.global_step,
               "filters": strategy.device_filters, "eval": res[0] if res else None},
              open(os.path.join(out_dir, f"result_{role.type}{role.index}.json"), "w"))


if __name__ == "__main__":
    {"mwms": mwms, "ps": ps_job}[sys.argv[1]](sys.argv[2])
