def complement(fn):
    return lambda *a, **k: not fn(*a, **k)


def compose(*fns):
    def run(x):
        for f in reversed(fns):
            x = f(x)
        return x

    return run


def identity(x):
    return x
