from . import functoolz, itertoolz  # noqa: F401
from .functoolz import *  # noqa: F401,F403
from .itertoolz import *  # noqa: F401,F403
