"""starframe_b200 — B200-native physics substep loop behind starframe's Physics/Body/Collider
surface.  The compute path is starframe_b200/csrc/libsfp_b200.so (hand-written sm_100a CUDA
behind the C ABI of include/sfp.h); importing the physics API without that library raises."""
from . import scenes  # noqa: F401
from ._ffi import (Api, Batch, BatchMulti, SfpError, World, BODY_DESC_DTYPE, BODY_STATE_DTYPE,  # noqa: F401
                   JOINT_DESC_DTYPE, PARAMS_DTYPE, PAIR_DTYPE, CONTACT_DTYPE, STATS_DTYPE,
                   FLAG_BATCH_WHOLE_WORLDS)
from .physics import (Body, Collider, Material, Physics, product_api, LIB_PATH)  # noqa: F401

__all__ = ["Physics", "Body", "Collider", "Material", "World", "Batch", "BatchMulti", "Api", "SfpError",
           "product_api", "scenes", "LIB_PATH"]
