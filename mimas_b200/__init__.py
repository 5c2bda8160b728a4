"""mimas_b200 -- B200 (sm_100a) CUDA implementation of the mimas-meshgen chunk mesher.

The package holds only what the hot path needs: the CUDA library (csrc/ ->
libmimas_meshgen_cuda.so, C ABI in include/mimas_meshgen_cuda.h) and the host-side mirror of
the reference crate API (meshgen.py) plus its block-table feeder (game_params.py).
"""
from .game_params import GameParams, load_params_from_toml  # noqa: F401
from .meshgen import (BatchMesh, BlockTextureIds, Blocky, ChunkMesh, CrossedStyle, HeadlessAssets,  # noqa: F401
                      Mesher, TextureIdCache, TextureIdObtainer, VERTEX_DTYPE, default_cache, gen_terrain_host, mesh_for_chunk)
