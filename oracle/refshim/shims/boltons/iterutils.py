def is_collection(obj):
    if isinstance(obj, (str, bytes, bytearray)):
        return False
    try:
        iter(obj)
        return True
    except TypeError:
        return False


def flatten_iter(iterable):
    for item in iterable:
        if is_collection(item) and not isinstance(item, dict):
            yield from flatten_iter(item)
        else:
            yield item


def flatten(iterable):
    return list(flatten_iter(iterable))
