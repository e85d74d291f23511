"""B200-native bundle adjustment + RANSAC hot path.

The directory name follows the build contract and is not a Python identifier; import it with
`importlib.import_module("extrinsic-camera-calibration-from-a-moving-person_b200")` or through
the top-level alias module `calib_b200`.
"""
