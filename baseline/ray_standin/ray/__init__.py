"""Minimal stand-in for the parts of Ray the reference imports. Stock torch only; see baseline/README.md."""
from . import train  # noqa: F401
from . import data  # noqa: F401

__version__ = "2.39.0-standin"
