"""b200ba — B200-native LM bundle adjustment and batched single-pose refinement.

The directory name carries a hyphen (the contract's package name), so import it through the
`b200ba` alias module at the repo root: `import b200ba`.
"""
from . import model_io, synth  # noqa: F401
from .api import (B200BundleAdjustment, B200PoseOptimizer, BAKernels, BAProblem, LMState,  # noqa: F401
                  LossFunction, launch_count, pose_mat_to_vec, rot_vec_to_mat, shard_points, shard_rows, unpack_params, with_noise)
from ._lib import B200BAError, LIB_PATH, SYMBOLS, load  # noqa: F401
