"""Helpers shared by the tests: build a product emitter for a tests/common.py configuration such that its
inputs are bit-identical to what the golden generator fed the reference."""
import numpy as np

import common


def golden(name):
    import os
    return np.load(os.path.join(common.GOLDEN_DIR, name + ".npz"))


def make(ssk, name, device=0, g=None, **override):
    spectral, kw = common.CONFIGS[name]
    kw = dict(kw)
    kw.update(override)
    normalized = False
    if "sun_direction" in kw and g is not None:
        # the reference normalises the property with an SSE rsqrt path that is CPU specific; hand our
        # library the exact unit vector the reference ended up with
        kw["sun_direction"] = [float(v) for v in g["state_sun_dir_world"]]
        normalized = True
    return ssk.SunskyEmitter(spectral=spectral, device=device, sun_direction_normalized=normalized, **kw)
