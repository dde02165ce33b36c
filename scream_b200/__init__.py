"""scream_b200 -- B200-native scene renderer for SCREAM's adversarial registration path.

Only what the hot path needs: the drop-in `RegistrationRender` module and the
batched `render_batch` op (scream_b200/render.py), their C-ABI CUDA library (scream_b200/csrc), the
item/point sharding helpers (scream_b200/shard.py) and synthetic clouds (scream_b200/synth.py).
"""
__all__ = ["RegistrationRender", "render_batch", "render_pair", "view_matrices", "make_offsets", "depth_range",
           "render_cloud_point_sharded"]


def __getattr__(name):
    if name in __all__:
        from . import render
        return getattr(render, name)
    raise AttributeError(name)
