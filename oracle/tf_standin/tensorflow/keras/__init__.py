"""TEST INFRASTRUCTURE ONLY (oracle): ``tensorflow.keras`` of the torch-backed stand-in (see ../__init__.py)."""
from . import backend  # noqa: F401
