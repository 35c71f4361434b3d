from .summary import SummaryWriter, read_scalars  # noqa: F401
from .tensorboard import start_tensorboard  # noqa: F401
from .profiler import ProfilerHook  # noqa: F401
from .logging_utils import set_verbosity, get_logger  # noqa: F401
