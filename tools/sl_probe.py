import sys, os, numpy as np
sys.path.insert(0, os.getcwd())
from cpp_volume_rendering_b200 import _capi, inputs
n, w, h = 512, 1024, 1024
vox = inputs.synth_volume(n, np.uint8)
table = inputs.tf_build_linear(inputs.bonsai_tf())
cam = inputs.default_camera(n, w, h); light = inputs.default_light(n)
ctx = _capi.Context(0)
ctx.upload_volume(vox); ctx.upload_tf(table)
r = _capi.Renderer(ctx, _capi.VRB_VCT)
r.init(w, h)
r.update(cam, light); r.redraw(); ctx.synchronize()
r.set_bricks(-1, 8, (2, 2, 2))
r.update(cam, light)
for i in range(2):
    r.redraw(); ctx.synchronize()
print("ok")
