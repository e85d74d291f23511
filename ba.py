#!/usr/bin/env python3
"""Drop-in for the reference's ba.py: same flags, same files in and out (results/<target>[_mask]_ba.json), same
Python API (to_theta / from_theta / objfun / ba_main) — computed by the sm_100a kernels of libbax.so."""
import calib_b200  # noqa: F401  (binds the hyphenated package directory)
from calib_b200.ba import OP_BONE, ba_main, from_theta, jacobian, objfun, to_theta  # noqa: F401
from calib_b200.cli import ba_cli, parse_args, save_mask  # noqa: F401

if __name__ == "__main__":
    ba_cli()
