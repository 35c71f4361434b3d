"""tf.estimator surface."""
from .estimator import (ModeKeys, RunConfig, ConfigProto, GPUOptions, TrainSpec, EvalSpec, FinalExporter,  # noqa: F401
                        LatestExporter, EstimatorSpec, Estimator, SessionRunHook, model_to_estimator,
                        train_and_evaluate, export, ServingInputReceiver, TensorServingInputReceiver,
                        session_config_from_env)
from .symbolic import placeholder, reduce_sum, reduce_mean, get_or_create_global_step  # noqa: F401
from ..obs.profiler import ProfilerHook  # noqa: F401
