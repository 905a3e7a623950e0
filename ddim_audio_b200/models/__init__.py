from .diffusion import Model  # noqa: F401
