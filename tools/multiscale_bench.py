"""Device time of the multiscaling redraws (render + filter pass) of the 1-pass ray caster against the plain redraw.
One JSON line per (mode, kernel).  usage: python tools/multiscale_bench.py [volume n] [screen]"""
import json
import sys

import numpy as np

sys.path.insert(0, ".")
from cpp_volume_rendering_b200 import _capi, inputs  # noqa: E402


def main():
    n = int(sys.argv[1]) if len(sys.argv) > 1 else 256
    s = int(sys.argv[2]) if len(sys.argv) > 2 else 1024
    ctx = _capi.Context(0)
    ctx.upload_volume(inputs.synth_volume(n, np.uint8)); ctx.upload_tf(inputs.tf_build_linear(inputs.bonsai_tf()))
    r = _capi.Renderer(ctx, _capi.VRB_RC1PASS)
    r.init(s, s)
    r.update(inputs.make_camera((n, n, 2.0 * n), width=s, height=s), inputs.default_light(256))

    def timed(mode, reps=20):
        for _ in range(3):
            r.redraw(mode)
        ms = []
        for _ in range(reps):
            r.redraw(mode); ms.append(r.counters()[1])
        return float(np.median(ms))

    base = timed(0)
    print(json.dumps({"mode": 0, "screen": s, "render": list(r.render_size()), "ms": round(base, 4)}))
    for mode, kernels in [(1, [1]), (2, range(6)), (3, range(6))]:
        for k in kernels:
            r.set_multiscaling(mode, k)
            plain = timed(0)
            full = timed(mode)
            print(json.dumps({"mode": mode, "kernel": k, "screen": s, "render": list(r.render_size()), "render_ms": round(plain, 4),
                              "redraw_ms": round(full, 4), "filter_ms": round(full - plain, 4)}))
    r.close(); ctx.close()


if __name__ == "__main__":
    main()
