"""deeplearning4j_b200 -- B200-native (sm_100a) back-end for Deeplearning4j's CNN layer helpers.

Only the hot path lives here: the C ABI (include/dl4jb200.h, csrc/), the host-side mirror of the four
helper classes the reference's layers look up by reflection (helpers.py), the device-resident step driver
for the BASELINE configs (network.py, zoo.py) and the data-parallel wrapper (parallel.py).
Importing the package does not need a GPU; creating a Context does. There is no CPU fallback.
"""
from . import _lib  # noqa: F401
from ._lib import (DL4JB200Error, DL4JInvalidConfigException, DL4JInvalidInputException,  # noqa: F401
                   HelperNotSupported)

__all__ = ["_lib", "helpers", "DL4JB200Error", "DL4JInvalidConfigException", "DL4JInvalidInputException",
           "HelperNotSupported"]
