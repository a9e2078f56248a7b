"""gpflow/features.py (1.5.1): the base classes gpsig/inducing_variables.py derives from; Kuu / Kuf are the dispatchers."""
from .dispatch import dispatch  # noqa: F401
from .params import Parameter


class InducingFeature:
    pass


class InducingPointsBase(InducingFeature):
    def __init__(self, Z):
        self.Z = Parameter(Z)

    def __len__(self):
        return self.Z.shape[0]


def Kuu(*args, **kwargs):
    raise NotImplementedError


def Kuf(*args, **kwargs):
    raise NotImplementedError
