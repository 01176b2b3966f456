class Template:
    def __init__(self, *a, **k):
        pass

    def render(self, *a, **k):
        raise NotImplementedError("mako shim: only the numpy backend is usable")
