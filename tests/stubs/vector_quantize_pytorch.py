"""Import stub (tests only): quantize.py imports these names; they are constructed only when the model is built with
quantize=True, which the render-path tests never do."""


class VectorQuantize:
    def __init__(self, *a, **k):
        raise NotImplementedError("vector_quantize_pytorch is not installed in this image (stub for import only)")


class ResidualVQ(VectorQuantize):
    pass
