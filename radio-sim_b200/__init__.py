"""radio-sim_b200 -- B200-native radio-medium hot path for joakimeriksson/radio-sim.

Only what the path needs lives here: `csrc/` (sm_100a CUDA kernels + the C-ABI of
include/radiomedium_b200.h, built into librmb200.so) and the host-side mirror of the
reference's RadioMedium interface.  Import as `radio_sim_b200` (a directory name with a
hyphen is not importable; radio_sim_b200/ is a shim pointing here).
"""
from ._abi import (MODEL_NULL, MODEL_UDGM, MODEL_UDGM_CONSTANT, MODEL_N2N, MODEL_SINR,
                   NOT_HEARD, INTERFERED, DELIVERED, LIB_PATH, SYMBOLS, load)
from .medium import Medium, MediumError, TickResult, Group, GroupResult

__all__ = ["Medium", "MediumError", "TickResult", "Group", "GroupResult", "load", "LIB_PATH", "SYMBOLS",
           "MODEL_NULL", "MODEL_UDGM", "MODEL_UDGM_CONSTANT", "MODEL_N2N", "MODEL_SINR",
           "NOT_HEARD", "INTERFERED", "DELIVERED"]
