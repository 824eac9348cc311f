from . import bbox, image  # noqa: F401
