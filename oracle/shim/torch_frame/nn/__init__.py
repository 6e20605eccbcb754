from . import encoder  # noqa: F401
