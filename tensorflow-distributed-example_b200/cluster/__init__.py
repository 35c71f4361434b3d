from .tf_config import *  # noqa: F401,F403
from .tf_config import (ClusterResolver, ClusterSpec, parse_tf_config, resolve_from_env,  # noqa: F401
                        synthesize_tf_config_from_cluster_spec, device_filters_from_tf_config,
                        make_tf_config)
