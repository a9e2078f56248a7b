def params_as_tensors(method):
    return method


class params_as_tensors_for:
    def __init__(self, *objs, **kwargs):
        pass

    def __enter__(self):
        return self

    def __exit__(self, *exc):
        return False


def autoflow(*specs):
    """Eager: the wrapped method is called on arrays directly."""
    def deco(method):
        return method
    return deco


def name_scope(name=None):
    def deco(method):
        return method
    return deco
