"""Kernel objects and readers that feed the device-resident feature pipeline."""
from . import io, kernels  # noqa: F401
