"""Importable alias for the package directory ``visual-inertial-odometry-msckf-stereo_b200/``
(whose name, fixed by the build contract, is not a valid Python identifier)."""
import os as _os

__path__ = [_os.path.join(_os.path.dirname(_os.path.dirname(_os.path.abspath(__file__))),
                          'visual-inertial-odometry-msckf-stereo_b200')]
