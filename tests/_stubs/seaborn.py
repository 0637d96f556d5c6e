def set(*a, **k):
    pass
