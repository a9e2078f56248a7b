"""``gpflow.dispatch.dispatch`` (multipledispatch): functions of one name registered per argument types; the call goes to
the first registration whose types match the positional arguments."""
_REGISTRY = {}


class _Dispatcher:
    def __init__(self, name):
        self.name, self.impls = name, []

    def __call__(self, *args, **kwargs):
        for types, fn in self.impls:
            if len(types) == len(args) and all(isinstance(a, t) for a, t in zip(args, types)):
                return fn(*args, **kwargs)
        raise NotImplementedError('no %s for %r' % (self.name, tuple(type(a).__name__ for a in args)))


def dispatch(*types):
    def deco(fn):
        d = _REGISTRY.setdefault(fn.__name__, _Dispatcher(fn.__name__))
        d.impls.append((types, fn))
        return d
    return deco
