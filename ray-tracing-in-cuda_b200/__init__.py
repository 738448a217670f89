"""B200-native path tracer — Python host mirror of the reference's render path.

The directory name contains hyphens (it follows the reference repository's name), so import it through
`__graft_entry__.load_package()`, which registers it as the module `rt_b200`.
"""
from . import capi, multi_gpu  # noqa: F401
from .renderer import Renderer, RtError, parse_scene, render_scene_file  # noqa: F401
