"""mpc_interface_b200 -- B200-native QP assembly behind the mpc-interface API.

Host mirror of the reference's formulation API (dynamics, combinations, goal,
restrictions, body, tools) whose matrix generation runs on hand-written
sm_100a CUDA kernels through the C-ABI library ``csrc/libmpci.so``
(include/mpci.h).  There is no CPU fallback: without the library and a GPU
every matrix-producing call raises.
"""
from . import tools  # noqa: F401
from .params import Param  # noqa: F401
from .combinations import LineCombo  # noqa: F401
from .dynamics import ControlSystem, ExtendedSystem, DomainVariable  # noqa: F401
from .goal import Cost  # noqa: F401
from .restrictions import Constraint, Box, box_boundaries  # noqa: F401
from .body import Formulation  # noqa: F401

__version__ = "0.1.0"
