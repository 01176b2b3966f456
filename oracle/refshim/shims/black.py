import enum


class TargetVersion(enum.Enum):
    PY36 = 6
    PY37 = 7
    PY38 = 8
    PY39 = 9
    PY310 = 10
    PY311 = 11
    PY312 = 12


class FileMode:
    def __init__(self, *a, **k):
        pass


Mode = FileMode


def format_str(src, *, mode=None):
    return src
