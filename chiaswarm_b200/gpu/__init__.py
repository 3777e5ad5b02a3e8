from .device import Device  # noqa: F401
