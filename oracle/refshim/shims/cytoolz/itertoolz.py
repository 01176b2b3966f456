import itertools


def unique(seq, key=None):
    seen = set()
    for x in seq:
        k = x if key is None else key(x)
        if k not in seen:
            seen.add(k)
            yield x


def _keyfn(key):
    if callable(key):
        return key
    if isinstance(key, list):
        return lambda x: tuple(x[k] for k in key)
    return lambda x: x[key]


def groupby(key, seq):
    key = _keyfn(key)
    out = {}
    for x in seq:
        out.setdefault(key(x), []).append(x)
    return out


_no = object()


def reduceby(key, binop, seq, init=_no):
    key = _keyfn(key)
    out = {}
    for x in seq:
        k = key(x)
        if k not in out:
            if init is _no:
                out[k] = x
                continue
            out[k] = init() if callable(init) else init
        out[k] = binop(out[k], x)
    return out


def partition(n, seq, pad=_no):
    args = [iter(seq)] * n
    if pad is _no:
        return zip(*args)
    return itertools.zip_longest(*args, fillvalue=pad)


def partition_all(n, seq):
    it = iter(seq)
    while True:
        chunk = tuple(itertools.islice(it, n))
        if not chunk:
            return
        yield chunk


def pluck(ind, seqs, default=_no):
    if isinstance(ind, list):
        return (tuple(s[i] for i in ind) for s in seqs)
    return (s[ind] for s in seqs)


def take_nth(n, seq):
    return itertools.islice(seq, 0, None, n)


def diff(*seqs, **kwargs):
    default = kwargs.get("default", _no)
    key = kwargs.get("key")
    if len(seqs) == 1 and not isinstance(seqs[0], (str, bytes)):
        seqs = seqs[0]
    it = zip(*seqs) if default is _no else itertools.zip_longest(*seqs, fillvalue=default)
    for items in it:
        ks = items if key is None else tuple(map(key, items))
        if any(k != ks[0] for k in ks[1:]):
            yield items


def first(seq):
    return next(iter(seq))


def concat(seqs):
    return itertools.chain.from_iterable(seqs)
