class classproperty:
    def __init__(self, fn):
        self.fn = fn

    def __get__(self, obj, objtype=None):
        return self.fn(objtype if objtype is not None else type(obj))


def make_sentinel(name="_MISSING", var_name=None):
    class Sentinel:
        def __init__(self):
            self.name = name
            self.var_name = var_name

        def __repr__(self):
            return self.var_name or "%s(%r)" % (type(self).__name__, self.name)

        def __bool__(self):
            return False

        def __copy__(self):
            return self

        def __deepcopy__(self, memo):
            return self

    return Sentinel()
