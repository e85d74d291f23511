#!/usr/bin/env python3
"""Drop-in for the reference's calib_linear.py: same flags, same files in and out (results/linear_<a>_<b>[_mask].json),
same Python API (calib_linear / main_linear) — computed by the sm_100a kernels of libbax.so."""
import calib_b200  # noqa: F401
from calib_b200.ba import OP_BONE  # noqa: F401
from calib_b200.calib_linear import calib_linear, linear_cli, main_linear  # noqa: F401
from calib_b200.cli import parse_args  # noqa: F401

if __name__ == "__main__":
    linear_cli()
