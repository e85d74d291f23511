#!/usr/bin/env python3
"""Drop-in for the reference's calib_ransac.py: same flags, same files in and out (results/ransac_<a>_<b>.json),
same Python API (calib_linear / calib_linear_ransac / main_ransac) — computed by the sm_100a kernels of libbax.so."""
import calib_b200  # noqa: F401
from calib_b200.ba import OP_BONE  # noqa: F401
from calib_b200.calib_ransac import calib_linear, calib_linear_ransac  # noqa: F401
from calib_b200.cli import main_ransac, parse_args, ransac_cli  # noqa: F401

if __name__ == "__main__":
    ransac_cli()
