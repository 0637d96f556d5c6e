def summary(*a, **k):
    return None
