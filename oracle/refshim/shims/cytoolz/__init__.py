from toolz import functoolz, itertoolz  # noqa: F401
from toolz.functoolz import *  # noqa: F401,F403
from toolz.itertoolz import *  # noqa: F401,F403
