#!/usr/bin/env python3
"""Drop-in for the reference's eval.py: same flags, same files in (gt_subset/, results/<target>.json) and out
(results/eval/<target>.json), same eval_main signature — triangulation and reprojection on the sm_100a kernels."""
import calib_b200  # noqa: F401
from calib_b200.cli import parse_args  # noqa: F401
from calib_b200.eval import align, calc_repro_error, create_val_summary, eval_cli, eval_main  # noqa: F401

if __name__ == "__main__":
    eval_cli()
