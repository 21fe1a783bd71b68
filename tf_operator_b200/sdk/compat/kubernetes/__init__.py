"""Minimal ``kubernetes`` stand-in: see tf_operator_b200/sdk/compat/__init__.py."""
from . import client  # noqa: F401
