"""B200-native (sm_100a) per-pixel reflectance fitting path of Smartphone-based-RTI.

Import as ``rti_b200`` (the repository root carries an alias package because this
directory's name is not a Python identifier).
"""
from . import constants
from ._capi import RtiError, RtiLibraryMissing, RtiSingularError
from .analysis import (compute, compute_intensities, install, interpolate_intensities,
                       interpolate_shared_light, light_axis, prepare_images_data, read_from_file,
                       write_on_file)

__all__ = [
    "constants", "compute", "compute_intensities", "interpolate_intensities", "prepare_images_data",
    "install", "interpolate_shared_light", "light_axis", "read_from_file", "write_on_file", "RtiError", "RtiLibraryMissing",
    "RtiSingularError",
]
