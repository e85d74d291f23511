"""The reference's shared argument parser surface (argument.py:7-58), re-exported for scripts that import it."""
import calib_b200  # noqa: F401
from calib_b200.cli import parse_args as _parse


def parse_args(predefined_args=None):
    return _parse(predefined_args)
