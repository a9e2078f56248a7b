"""Transforms only constrain the optimiser's parametrisation; evaluated kernels see the constrained value."""
positive = 'positive'


class Logistic:
    def __init__(self, a=0., b=1.):
        self.a, self.b = a, b
