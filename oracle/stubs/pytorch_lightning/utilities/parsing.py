import inspect


def get_init_args(frame):
    """Named ``__init__`` parameters of the calling frame, merged with its ``**kwargs`` dict."""
    local_vars = frame.f_locals
    if '__class__' not in local_vars:
        return {}
    cls = local_vars['__class__']
    sig = inspect.signature(cls.__init__)
    out = {}
    for name, p in sig.parameters.items():
        if name == 'self':
            continue
        if p.kind == inspect.Parameter.VAR_POSITIONAL:
            continue
        if p.kind == inspect.Parameter.VAR_KEYWORD:
            out.update(local_vars.get(name, {}))
            continue
        if name in local_vars:
            out[name] = local_vars[name]
    return out
