class VtkGroup:
    def __init__(self, *_a, **_k):
        raise NotImplementedError
