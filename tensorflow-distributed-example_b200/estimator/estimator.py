"""Estimator-like training driver: RunConfig / TrainSpec / EvalSpec / FinalExporter /
``model_to_estimator`` / ``train_and_evaluate``.

Reference call sites: ``RunConfig(experimental_distribute=DistributeConfig(train_distribute=PS,
eval_distribute=Mirrored), session_config, model_dir, save_summary_steps=100, log_step_count_steps=100,
save_checkpoints_steps=500)`` (mnist_keras_distributed.py:240-248), ``RunConfig(train_distribute=...)``
(tf2_mnist_distributed.py:205), ``model_to_estimator(keras_model, model_dir, config)``
(mnist_keras_distributed.py:118-119), ``TrainSpec(input_fn, max_steps[, hooks])`` (:255-262),
``FinalExporter('exporter', serving_input_fn)`` (:264), ``EvalSpec(input_fn, steps=None, name,
exporters, start_delay_secs=10, throttle_secs=10)`` (:265-275), ``train_and_evaluate`` (:283).

What TensorFlow did behind those calls and this module does itself: dispatch on
the TF_CONFIG task type (ps -> serve variables; chief/master -> train + chief-only checkpoints and
summaries; worker -> train; evaluator -> evaluate every new checkpoint honouring start_delay/throttle),
auto-resume from the latest checkpoint in ``model_dir``, keep the last 5 checkpoints, write ``loss`` and
``global_step/sec`` scalars, log "loss = ..., step = ..." lines, and run the exporters after the final
evaluation. The hot loop itself is ``Sequential.train_on_batch`` (one CUDA-graph launch per step); hooks
run on the host around it and only force a device sync at their own cadence.
"""
from __future__ import annotations

import logging
import math
import os
import time
from typing import Callable, Dict, List, Optional

import numpy as np
import torch

from ..ckpt import checkpoint as ckpt
from ..cluster import tf_config as tfc
from ..obs.summary import SummaryWriter
from ..parallel.strategy import Strategy, get_strategy
from . import symbolic as sym

log = logging.getLogger("tfde_b200")


class ModeKeys:
    TRAIN = "train"
    EVAL = "eval"
    PREDICT = "infer"


class RunConfig:
    def __init__(self, model_dir=None, tf_random_seed=None, save_summary_steps=100,
                 save_checkpoints_steps=None, save_checkpoints_secs=None, session_config=None,
                 keep_checkpoint_max=5, log_step_count_steps=100, train_distribute=None,
                 eval_distribute=None, experimental_distribute=None, device_fn=None):
        self.model_dir = model_dir
        self.tf_random_seed = tf_random_seed
        self.save_summary_steps = save_summary_steps
        if save_checkpoints_steps is None and save_checkpoints_secs is None:
            save_checkpoints_secs = 600   # Estimator default (tf2_mnist_distributed.py relies on it)
        self.save_checkpoints_steps = save_checkpoints_steps
        self.save_checkpoints_secs = save_checkpoints_secs
        self.session_config = session_config
        self.keep_checkpoint_max = keep_checkpoint_max
        self.log_step_count_steps = log_step_count_steps
        if experimental_distribute is not None:
            train_distribute = train_distribute or experimental_distribute.train_distribute
            eval_distribute = eval_distribute or experimental_distribute.eval_distribute
        self.train_distribute = train_distribute
        self.eval_distribute = eval_distribute
        self.cluster = tfc.resolve_from_env()

    @property
    def is_chief(self) -> bool:
        return self.cluster.is_chief

    @property
    def task_type(self) -> str:
        return self.cluster.task.type

    @property
    def task_id(self) -> int:
        return self.cluster.task.index

    def replace(self, **kw):
        for k, v in kw.items():
            setattr(self, k, v)
        return self


class ConfigProto:
    """tf.ConfigProto(device_filters=[...], gpu_options=...) as carried by RunConfig.session_config
    (mnist_keras_distributed.py:176-189)."""

    def __init__(self, device_filters=None, gpu_options=None):
        self.device_filters = list(device_filters) if device_filters else []
        self.gpu_options = gpu_options
        if gpu_options is not None and getattr(gpu_options, "per_process_gpu_memory_fraction", None):
            if torch.cuda.is_available():
                torch.cuda.set_per_process_memory_fraction(gpu_options.per_process_gpu_memory_fraction)


class GPUOptions:
    def __init__(self, per_process_gpu_memory_fraction=None):
        self.per_process_gpu_memory_fraction = per_process_gpu_memory_fraction


def session_config_from_env(environ=None) -> Optional[ConfigProto]:
    """``_get_session_config_from_env_var`` (mnist_keras_distributed.py:165-189)."""
    import json
    env = os.environ if environ is None else environ
    cfg = json.loads(env.get("TF_CONFIG", "{}"))
    f = tfc.device_filters_from_tf_config(cfg)
    return ConfigProto(device_filters=f) if f else None


class TrainSpec:
    def __init__(self, input_fn: Callable, max_steps=None, hooks=None):
        self.input_fn, self.max_steps, self.hooks = input_fn, max_steps, list(hooks or [])


class EvalSpec:
    def __init__(self, input_fn: Callable, steps=100, name=None, hooks=None, exporters=None,
                 start_delay_secs=120, throttle_secs=600):
        self.input_fn, self.steps, self.name = input_fn, steps, name
        self.hooks = list(hooks or [])
        if exporters is None:
            exporters = []
        elif not isinstance(exporters, (list, tuple)):
            exporters = [exporters]
        self.exporters = list(exporters)
        self.start_delay_secs, self.throttle_secs = start_delay_secs, throttle_secs


class ServingInputReceiver:
    def __init__(self, features, receiver_tensors):
        self.features, self.receiver_tensors = features, receiver_tensors

    def signature(self) -> Dict:
        t = self.receiver_tensors
        if isinstance(t, dict):
            t = next(iter(t.values()))
        return {"name": getattr(t, "name", "input"), "shape": getattr(t, "shape", None),
                "dtype": getattr(t, "dtype", "float32")}


class TensorServingInputReceiver(ServingInputReceiver):
    """tf.estimator.export.TensorServingInputReceiver(features, placeholder)
    (mnist_keras_distributed.py:161-162)."""


class export:
    ServingInputReceiver = ServingInputReceiver
    TensorServingInputReceiver = TensorServingInputReceiver


class FinalExporter:
    """Exports once, after the last evaluation, to ``<model_dir>/export/<name>/<timestamp>/``."""

    def __init__(self, name, serving_input_receiver_fn):
        self.name, self.serving_input_receiver_fn = name, serving_input_receiver_fn

    def export(self, estimator, export_path, checkpoint_path, eval_result, is_the_final_export):
        if not is_the_final_export:
            return None
        return estimator.export_saved_model(export_path, self.serving_input_receiver_fn)


class LatestExporter(FinalExporter):
    def export(self, estimator, export_path, checkpoint_path, eval_result, is_the_final_export):
        return estimator.export_saved_model(export_path, self.serving_input_receiver_fn)


class EstimatorSpec:
    def __init__(self, mode=None, predictions=None, loss=None, train_op=None, eval_metric_ops=None,
                 labels=None, **_):
        self.mode, self.predictions, self.loss, self.train_op = mode, predictions, loss, train_op
        self.eval_metric_ops = eval_metric_ops


# ----------------------------------------------------------------------------------------------
# hooks (SessionRunHook protocol reduced to what a static step needs)
# ----------------------------------------------------------------------------------------------
class SessionRunHook:
    def begin(self, estimator=None): ...
    def before_run(self, step, model=None): ...
    def after_run(self, step, model=None, logs=None): ...
    def end(self, estimator=None): ...


class _LoggingHook(SessionRunHook):
    """StepCounterHook + SummarySaverHook + the 'loss = ..., step = ...' line, chief only for files."""

    def __init__(self, config: RunConfig, writer: Optional[SummaryWriter], batch_fn):
        self.cfg, self.w, self.batch_fn = config, writer, batch_fn
        self.t_last, self.s_last = None, None
        self.rates: List[float] = []

    def begin(self, estimator=None):
        self.t_last = None

    def after_run(self, step, model=None, logs=None):
        done = step + 1
        every_log = self.cfg.log_step_count_steps
        every_sum = self.cfg.save_summary_steps
        want_log = every_log and done % every_log == 0
        want_sum = every_sum and done % every_sum == 0
        if not (want_log or want_sum or done == 1):
            return
        loss = model.last_step_loss()     # the only device sync, at hook cadence
        now = time.time()
        if self.w is not None and (want_sum or done == 1):
            self.w.scalar("loss", loss, done)
        if want_log:
            if self.t_last is not None and now > self.t_last:
                rate = (done - self.s_last) / (now - self.t_last)
                self.rates.append(rate)
                log.info("global_step/sec: %g", rate)
                if self.w is not None:
                    self.w.scalar("global_step/sec", rate, done)
                    self.w.scalar("images/sec", rate * self.batch_fn(), done)
            self.t_last, self.s_last = now, done
        if want_log or done == 1:
            log.info("loss = %s, step = %d", loss, done)
        if self.w is not None:
            self.w.flush()


# ----------------------------------------------------------------------------------------------
# Estimator
# ----------------------------------------------------------------------------------------------
class Estimator:
    def __init__(self, model_fn=None, model_dir=None, config: Optional[RunConfig] = None, params=None,
                 keras_model=None, keras_model_builder=None):
        self.config = config or RunConfig(model_dir=model_dir)
        self.model_dir = model_dir or self.config.model_dir
        if self.model_dir is None:
            import tempfile
            self.model_dir = tempfile.mkdtemp(prefix="tfde_estimator_")
        self.config.model_dir = self.model_dir
        self.model_fn, self.params = model_fn, params
        self._keras_model = keras_model
        self._builder = keras_model_builder
        self._models: Dict[str, object] = {}
        self._train_global_batch = None
        self.global_step = 0
        self.last_eval: Optional[Dict] = None

    # ---- model plumbing -----------------------------------------------------------------------
    def _strategy_for(self, mode) -> Strategy:
        st = self.config.train_distribute if mode == ModeKeys.TRAIN else (
            self.config.eval_distribute or self.config.train_distribute)
        return st if st is not None else get_strategy()

    def _model(self, mode):
        """One model instance per (strategy) -- train and eval share it when they share the strategy's
        device; eval under a different strategy gets the weights through the checkpoint, exactly the
        channel TF uses between training and evaluation jobs."""
        st = self._strategy_for(mode)
        key = id(st)
        if key in self._models:
            return self._models[key]
        if self._keras_model is not None and (not self._models) and self._keras_model.strategy is st:
            model = self._keras_model
        elif self._keras_model is not None and not self._models and self._keras_model.device == st.device \
                and st.num_replicas_in_sync == 1 and self._keras_model.strategy.num_replicas_in_sync == 1:
            model = self._keras_model
        elif self.model_fn is not None:
            with st.scope():
                model = self._build_from_model_fn(mode)
        else:
            with st.scope():
                model = self._clone_keras_model()
        self._models[key] = model
        return model

    def _clone_keras_model(self):
        from ..keras import layers as L
        from ..keras.models import Sequential
        import copy
        src = self._keras_model
        if self._builder is not None:
            return self._builder()
        if src is None:
            raise ValueError("Estimator needs a keras_model or a model_fn")
        m = Sequential([L.layer_from_config(type(l).__name__, l.get_config()) for l in src.layers],
                       name=src.name, seed=src.seed)
        if src.compiled:
            m.compile(optimizer=copy.copy(src.optimizer), loss=src.loss,
                      metrics=[n for n in src.metrics_names if n != "loss"])
            m.optimizer._bound_to = None
            m.optimizer.hparams = None
        if not m.built and src.built:
            m.build(src.input_shape)
        if src.built and m.strategy.num_replicas_in_sync == 1 and src.device == m.device:
            m.set_weights(src.get_weights())
        return m

    def _build_from_model_fn(self, mode):
        features = sym.SymbolicTensor("features")
        labels = sym.SymbolicTensor("labels", dtype="int64")
        spec = self.model_fn(features, labels, mode)
        if not isinstance(spec, EstimatorSpec):
            raise TypeError("model_fn must return an EstimatorSpec")
        self._spec = spec
        if mode == ModeKeys.TRAIN:
            op = spec.train_op
            if not isinstance(op, sym.TrainOp):
                raise TypeError("EstimatorSpec.train_op must come from optimizer.minimize(loss, global_step)")
            model = op.model
            if not model.compiled:
                model.compile(optimizer=op.optimizer, loss=op.loss.loss_obj, metrics=["accuracy"])
            self._train_global_batch = op.loss.global_batch
            return model
        loss = spec.loss
        if isinstance(loss, sym.SymbolicLoss):
            model = loss.model
            if not model.compiled:
                from ..keras import optimizers
                model.compile(optimizer=optimizers.SGD(0.0), loss=loss.loss_obj, metrics=["accuracy"])
            return model
        pred = spec.predictions
        if isinstance(pred, dict):
            pred = next(iter(pred.values()))
        if isinstance(pred, sym.SymbolicOutput):
            return pred.model
        raise TypeError("cannot find the model behind the EstimatorSpec")

    # ---- checkpoints --------------------------------------------------------------------------
    def latest_checkpoint(self) -> Optional[str]:
        return ckpt.latest_checkpoint(self.model_dir)

    def _restore_latest(self, model, need_optimizer=True) -> int:
        p = self.latest_checkpoint()
        if p is None:
            return 0
        if not model.built:
            model.build()
        step = ckpt.restore_checkpoint(model, p)
        log.info("Restoring parameters from %s (global_step %d)", p, step)
        return step

    # ---- train ----------------------------------------------------------------------------------
    def train(self, input_fn, hooks=None, steps=None, max_steps=None, saving_listeners=None):
        cfg = self.config
        model = self._model(ModeKeys.TRAIN)
        st = model.strategy
        if not model.built:
            model.build()
        model._bind_optimizer()
        self.global_step = self._restore_latest(model)
        if hasattr(st, "on_train_begin"):
            st.on_train_begin(model)
            if hasattr(st, "global_step"):
                self.global_step = max(self.global_step, st.global_step())
        target = math.inf
        if max_steps is not None:
            target = max_steps          # may be a float (468.75 in the reference): loop while step < target
        if steps is not None:
            target = min(target, self.global_step + steps)
        if self.global_step >= target:
            log.info("Skipping training since max_steps has already been reached.")
            return self
        chief = cfg.is_chief and st.is_chief
        if chief:
            os.makedirs(self.model_dir, exist_ok=True)
        writer = SummaryWriter(self.model_dir) if chief else None
        n_rep = st.num_replicas_in_sync if st.synchronous else 1
        last_batch = [0]
        all_hooks = [_LoggingHook(cfg, writer, lambda: last_batch[0] * n_rep)] + list(hooks or [])
        for h in all_hooks:
            h.begin(self)
        ds = input_fn()
        it = iter(ds)
        t_ckpt = time.time()
        last_saved = self.global_step if self.latest_checkpoint() else -1
        if chief and last_saved < 0:
            ckpt.save_checkpoint(model, self.model_dir, self.global_step, cfg.keep_checkpoint_max)
            last_saved = self.global_step
        elif not chief and getattr(model.arena, "state_sharded", False) and last_saved < 0:
            ckpt.save_checkpoint(model, self.model_dir, self.global_step, cfg.keep_checkpoint_max)
        while self.global_step < target:
            try:
                elem = next(it)
            except StopIteration:
                log.info("input_fn exhausted at step %d", self.global_step)
                break
            x, y = elem[0], elem[1]
            b = int(x.shape[0])
            last_batch[0] = b
            gb = self._train_global_batch or b * n_rep
            step = self.global_step
            for h in all_hooks:
                h.before_run(step, model)
            model.train_on_batch(_t(x), _t(y), global_batch=gb)
            for h in all_hooks:
                h.after_run(step, model)
            self.global_step = st.after_step(self.global_step) if hasattr(st, "after_step") else step + 1
            due = False
            if cfg.save_checkpoints_steps:
                due = self.global_step - max(last_saved, 0) >= cfg.save_checkpoints_steps
            elif cfg.save_checkpoints_secs:
                due = time.time() - t_ckpt >= cfg.save_checkpoints_secs
            if due and st.synchronous and st.num_replicas_in_sync > 1 and not cfg.save_checkpoints_steps:
                due = False   # wall-clock triggers are not replica-consistent; only step triggers are collective
            if due and (chief or getattr(model.arena, "state_sharded", False)):
                ckpt.save_checkpoint(model, self.model_dir, self.global_step, cfg.keep_checkpoint_max)
                last_saved, t_ckpt = self.global_step, time.time()
        model.sync()
        if hasattr(st, "on_train_end"):
            st.on_train_end(model)      # parameter-server mode: final pull, release the ps tasks
        if (chief or getattr(model.arena, "state_sharded", False)) and last_saved != self.global_step:
            ckpt.save_checkpoint(model, self.model_dir, self.global_step, cfg.keep_checkpoint_max)
        for h in all_hooks:
            h.end(self)
        if writer is not None:
            writer.close()
        log.info("Loss for final step: %s.", model.last_step_loss())
        return self

    # ---- evaluate / predict / export -------------------------------------------------------------
    def evaluate(self, input_fn, steps=None, hooks=None, checkpoint_path=None, name=None) -> Dict:
        model = self._model(ModeKeys.EVAL)
        if not model.built:
            model.build()
        p = checkpoint_path or self.latest_checkpoint()
        step = self.global_step
        train_model = self._models.get(id(self._strategy_for(ModeKeys.TRAIN)))
        if p is not None and model is not train_model:
            ckpt.load_weights(model, p, strict=True)
            step = ckpt.step_of(p)
        elif p is not None:
            step = max(step, ckpt.step_of(p)) if model is train_model else ckpt.step_of(p)
        res = model.evaluate(input_fn(), steps=steps, return_dict=True)
        res["global_step"] = step
        if self.config.is_chief and model.strategy.is_chief:
            d = os.path.join(self.model_dir, "eval" if not name else f"eval_{name}")
            w = SummaryWriter(d)
            for k, v in res.items():
                if k != "global_step":
                    w.scalar(k, v, step)
            w.close()
        log.info("Saving dict for global step %d: %s", step, res)
        self.last_eval = res
        return res

    def predict(self, input_fn, predict_keys=None):
        model = self._model(ModeKeys.PREDICT)
        p = self.latest_checkpoint()
        if p is not None:
            ckpt.load_weights(model, p)
        for elem in input_fn():
            x = elem[0] if isinstance(elem, tuple) else elem
            out = model.predict(x, batch_size=len(x))
            for row in out:
                yield {"logits": row} if predict_keys is None or "logits" in predict_keys else row

    def export_saved_model(self, export_dir_base, serving_input_receiver_fn=None, checkpoint_path=None):
        model = self._model(ModeKeys.PREDICT)
        if not model.built:
            model.build()
        p = checkpoint_path or self.latest_checkpoint()
        train_model = self._models.get(id(self._strategy_for(ModeKeys.TRAIN)))
        if p is not None and model is not train_model:
            ckpt.load_weights(model, p)
        rcv = serving_input_receiver_fn() if serving_input_receiver_fn is not None else None
        if not (self.config.is_chief and model.strategy.is_chief):
            return None
        return ckpt.export_saved_model(model, export_dir_base, rcv)


def _t(a):
    if isinstance(a, torch.Tensor):
        return a
    a = np.ascontiguousarray(a)
    if a.dtype == np.float64:
        a = a.astype(np.float32)
    return torch.from_numpy(a)


def model_to_estimator(keras_model=None, model_dir=None, config: Optional[RunConfig] = None,
                       keras_model_builder=None) -> Estimator:
    """tf.keras.estimator.model_to_estimator (mnist_keras_distributed.py:118-119). The compiled Keras
    model (loss, optimizer, metrics) becomes the Estimator's model; when the RunConfig carries a
    distribution strategy the model is re-instantiated under that strategy's scope."""
    if keras_model is not None and not keras_model.compiled:
        raise ValueError("compile() the Keras model before model_to_estimator")
    return Estimator(model_dir=model_dir, config=config, keras_model=keras_model,
                     keras_model_builder=keras_model_builder)


# ----------------------------------------------------------------------------------------------
# train_and_evaluate
# ----------------------------------------------------------------------------------------------
def _final_export(estimator, eval_spec, result, final=True):
    for ex in eval_spec.exporters:
        base = os.path.join(estimator.model_dir, "export", ex.name)
        ex.export(estimator, base, estimator.latest_checkpoint(), result, final)


def _run_evaluator(estimator: Estimator, train_spec: TrainSpec, eval_spec: EvalSpec):
    """Continuous evaluation job: wait start_delay_secs, then evaluate every NEW checkpoint at most every
    throttle_secs until the checkpoint at max_steps has been scored; then run the exporters."""
    time.sleep(max(0, eval_spec.start_delay_secs))
    seen, result = None, None
    idle_since = time.time()
    while True:
        p = estimator.latest_checkpoint()
        if p is not None and p != seen:
            t0 = time.time()
            result = estimator.evaluate(eval_spec.input_fn, steps=eval_spec.steps, checkpoint_path=p,
                                        name=eval_spec.name)
            seen, idle_since = p, time.time()
            if train_spec.max_steps is not None and result["global_step"] >= train_spec.max_steps:
                break
            wait = eval_spec.throttle_secs - (time.time() - t0)
            if wait > 0:
                time.sleep(wait)
        else:
            if time.time() - idle_since > float(os.environ.get("TFDE_EVALUATOR_IDLE_TIMEOUT", "3600")):
                log.warning("evaluator: no new checkpoint, giving up")
                break
            time.sleep(min(1.0, max(0.05, eval_spec.throttle_secs)))
    _final_export(estimator, eval_spec, result, True)
    return result


def train_and_evaluate(estimator: Estimator, train_spec: TrainSpec, eval_spec: EvalSpec):
    """tf.estimator.train_and_evaluate (mnist_keras_distributed.py:283, tf2_mnist_distributed.py:237-241).
    Returns (eval_result, export_results) like TF when running locally, None for non-evaluating tasks."""
    cfg = estimator.config
    cluster = cfg.cluster
    task = cluster.task.type if cluster.distributed else None
    train_st = cfg.train_distribute

    if task == "ps":
        if train_st is None or not hasattr(train_st, "run_server"):
            raise RuntimeError("a 'ps' task needs RunConfig(train_distribute=ParameterServerStrategy())")
        train_st.run_server()     # returns when the chief announces the end of training
        return None
    if task == "evaluator":
        return _run_evaluator(estimator, train_spec, eval_spec), None

    if not cluster.distributed:
        # local loop: train up to the next checkpoint boundary, evaluate (throttled), repeat
        max_steps = train_spec.max_steps if train_spec.max_steps is not None else math.inf
        chunk = cfg.save_checkpoints_steps or None
        t_last_eval, result = 0.0, None
        t_start = time.time()
        while True:
            before = estimator.global_step
            estimator.train(train_spec.input_fn, hooks=train_spec.hooks, steps=chunk, max_steps=train_spec.max_steps)
            done = estimator.global_step >= max_steps or estimator.global_step == before
            now = time.time()
            if done or (now - t_start >= eval_spec.start_delay_secs and now - t_last_eval >= eval_spec.throttle_secs):
                result = estimator.evaluate(eval_spec.input_fn, steps=eval_spec.steps, name=eval_spec.name)
                t_last_eval = time.time()
                if not done:
                    _final_export(estimator, eval_spec, result, False)
            if done:
                break
        _final_export(estimator, eval_spec, result, True)
        return result, None

    # distributed training task (chief / master / worker)
    estimator.train(train_spec.input_fn, hooks=train_spec.hooks, max_steps=train_spec.max_steps)
    has_evaluator = cluster.cluster.num_tasks("evaluator") > 0
    if cfg.is_chief and (task == "master" or not has_evaluator):
        result = estimator.evaluate(eval_spec.input_fn, steps=eval_spec.steps, name=eval_spec.name)
        _final_export(estimator, eval_spec, result, True)
        return result, None
    return None
