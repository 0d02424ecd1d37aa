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
    # The raw property goes in, exactly as the golden generator handed it to the reference: the library
    # normalises it with the reference's own op order (dr::normalize on a ScalarVector3f, sunsky.cpp:922).
    return ssk.SunskyEmitter(spectral=spectral, device=device, **kw)
