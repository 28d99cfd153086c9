"""voxel_game_b200 — B200-native (sm_100a CUDA) voxel hot path of Devon7925/voxel-game behind the
reference's compute-pass / ray-pass interface. See DESIGN.md and include/voxelb200.h.

(The package directory is `voxel_game_b200`: a hyphen is not importable in Python.)
"""
from ._lib import COLLISION_DTYPE, PLAYER_DTYPE, PROJECTILE_DTYPE, RASTER_INSTANCE_DTYPE, RayCaps, VxbError, load  # noqa: F401
from .pipeline import (  # noqa: F401
    CHUNK_SIZE,
    CHUNK_VOLUME,
    Direction,
    Game,
    GameSettings,
    GameState,
    GraphicsSettings,
    PointLightingSystem,
    RasterizerSystem,
    SlabGroup,
    SlabPipeline,
    VoxelComputePipeline,
    WorldGen,
    initial_start_pos,
    perspective,
    quat_to_mat4,
    translation,
)
from .replay import ReplayGame, read_replay  # noqa: F401
