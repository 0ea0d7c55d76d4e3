"""trajectorycrafter_b200 — sm_100a-native forward re-rendering (``Warper.forward_warp``).

Drop-in for the point-cloud render stage of TrajectoryCrafter (reference models/utils.py:213-682):
the same ``Warper`` class and pose helpers on the Python side, hand-written CUDA behind the C ABI of
``include/tcwarp.h`` (``libtcwarp.so``).  There is no CPU or torch-eager fallback: importing the
package is cheap, but constructing a ``Warper`` without the built CUDA library raises.
"""
from .poses import (generate_traj_specified, generate_traj_txt, get_poses, sphere2pose,  # noqa: F401
                    txt_interpolation)
from .matrices import prepare_matrices  # noqa: F401
from .warper import Warper  # noqa: F401
from .pointcloud import GlobalPointCloudWarper  # noqa: F401

__all__ = ["Warper", "GlobalPointCloudWarper", "prepare_matrices", "sphere2pose", "generate_traj_specified",
           "generate_traj_txt", "txt_interpolation", "get_poses"]
