"""GPU: the 1 M-triangle mesh at 4K, 256 spp, bounce 0 from the tile candidate lists vs from the tree (RT_RENDER_NO_TILE_LISTS): pixels that
differ and the ray counts (found the grazing-triangle acceptance bug of round 2: 3 pixels in 2.1 G paths)."""
import sys; sys.path.insert(0,'.')
import numpy as np
import __graft_entry__ as ge
rt=ge.load_package()
from rt_b200 import scenes, capi
r=rt.Renderer(0)
ts=scenes.mesh_scene(708,708,3840,2160,256)
r.upload(ts)
r.stats_reset(); r.render(seed=0x2023); a,_,sa=r.download()
r.stats_reset(); r.render(seed=0x2023,flags=capi.RT_RENDER_NO_TILE_LISTS); b,_,sb=r.download()
diff=(a!=b).any(-1)
print("differing pixels", diff.sum(), "rays", sa["rays"], sb["rays"])
ys,xs=np.nonzero(diff)
for y,x in list(zip(ys,xs))[:10]:
    print(x,y,a[y,x],b[y,x], "tile", (y//4)*(3840//8)+x//8)
# AOV first-hit compare not available for lists; print count by row band
print(np.bincount(ys//216, minlength=10))
