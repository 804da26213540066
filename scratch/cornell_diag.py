import numpy as np, sys
sys.path.insert(0,'.')
from ipu_path_trace_b200 import cornell
g=np.load('tests/golden/cornell.npz')
rgb,depth,st=cornell.trace_paths(cornell.CornellOptions(), g['cam_xy'], g['rng_state'])
bad=np.nonzero((depth!=g['depth_default'])|(st!=g['state_default']).any(axis=1)|(rgb.view(np.uint32)!=g['rgb_default'].view(np.uint32)).any(axis=1))[0]
print("mismatching paths:", len(bad), "of", len(depth))
for i in bad[:12]:
    print(i, "cam", g['cam_xy'][i], "depth", depth[i], g['depth_default'][i], "rgb", rgb[i], g['rgb_default'][i], "state eq", (st[i]==g['state_default'][i]).all())
# depth-0/1 only probes: straight rays at each object
