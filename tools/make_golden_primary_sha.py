#!/usr/bin/env python
"""sha256 of the ORACLE's primary-hit arrays (type, index, t bits) and of its radiance bits for one frame of every
BASELINE.json configuration at its real size -> tests/golden/primary_sha.json.  The full-size GPU tests compare
the CUDA paths with the oracle AND the oracle with these digests, so a drift of the oracle itself is visible
(the reference ships no golden vectors: SURVEY.md §8(c))."""
import hashlib, json, os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np
import bench_scenes
from oracle.binding import Oracle
from tests.util import bits

FRAME = 3


def digests(cfg):
    scene, cam, W, H, kw = bench_scenes.build(cfg)
    o = Oracle(W, H, **kw)
    o.write_scene(scene.buffers())
    o.write_uniforms(FRAME, 0, cam.viewMatrix)
    rad, t, i, d = o.probe()
    return {"frame": FRAME, "size": [W, H],
            "type": hashlib.sha256(np.ascontiguousarray(t, np.int32).tobytes()).hexdigest(),
            "index": hashlib.sha256(np.ascontiguousarray(i, np.int32).tobytes()).hexdigest(),
            "t": hashlib.sha256(bits(d).tobytes()).hexdigest(),
            "radiance": hashlib.sha256(bits(rad).tobytes()).hexdigest(),
            "triangle_pixels": int((t == 2).sum())}


if __name__ == "__main__":
    out = {}
    for cfg in ("C1", "C2", "C3", "C4", "C5"):
        t0 = time.time()
        out[cfg] = digests(cfg)
        print(cfg, "%.1f s" % (time.time() - t0), out[cfg]["triangle_pixels"], flush=True)
    json.dump(out, open(os.path.join(ROOT, "tests", "golden", "primary_sha.json"), "w"), indent=1, sort_keys=True)
