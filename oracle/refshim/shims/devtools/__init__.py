def debug(*a, **k):
    return a[0] if a else None
