def Fire(*args, **kwargs):
    return None
