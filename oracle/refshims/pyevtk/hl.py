def gridToVTK(*_a, **_k):
    raise NotImplementedError


def pointsToVTK(*_a, **_k):
    raise NotImplementedError
