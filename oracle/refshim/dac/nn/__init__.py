from . import layers, quantize  # noqa
