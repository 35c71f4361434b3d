"""placeholder (filled in below in this build)."""
from .strategy import Strategy


class ParameterServerStrategy(Strategy):
    pass
