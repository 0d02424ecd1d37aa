"""ssk_update_batch_noalloc (include/sunsky_b200.h): parameters_changed with the setup quadrature's scratch in a
file-scope `static __device__` array instead of cudaMalloc'ed buffers. No counterpart in the reference (operators'
request). Checked against the existing entry point, ssk_update_batch, through the C ABI: after the same parameters
both leave a context in the same state, bit for bit (the whole ssk_state and the radiance / samples computed from
it) -- the tolerance every setup test here uses -- on two calls in a row, the second with different parameters so
that anything left in the shared array by the first call would show.

Collected last on purpose (`test_zz_`)."""
import numpy as np
import pytest

import common

pytestmark = pytest.mark.gpu


def _params(e, props):
    """the ssk_params `e.update(**props)` would pass"""
    from sunsky_b200._capi import Params
    p = Params.from_buffer_copy(e._params)
    e._fill(p, props, creating=False)
    return p


def _call(L, fn, emitters, props_list):
    from sunsky_b200 import _capi
    n = len(emitters)
    arr = (_capi.Params * n)(*[_params(e, p) for e, p in zip(emitters, props_list)])
    ctxs = (_capi.CTX * n)(*[e._ctx for e in emitters])
    rc = getattr(L, fn)(ctxs, arr, n)
    if rc == 0:
        for e, p in zip(emitters, arr):
            e._params = _capi.Params.from_buffer_copy(p)
    return rc


@pytest.mark.parametrize("spectral", [False, True])
def test_noalloc_update_equals_update_batch_twice_in_a_row(ssk, spectral):
    from sunsky_b200 import _capi
    import sunsky_b200.sweep as sw
    L = _capi.lib()
    n = 2                                                       # SSK_NOALLOC_MAX_EMITTERS
    base = dict(turbidity=3.0, albedo=0.3, sun_direction=sw.sun_direction(30.0))
    a = [ssk.SunskyEmitter(spectral=spectral, **base) for _ in range(n)]      # new entry point
    b = [ssk.SunskyEmitter(spectral=spectral, **base) for _ in range(n)]      # existing entry point
    rng = np.random.default_rng(11)
    wo = common.hemisphere_dirs(513, rng); u = rng.random((2, 513), dtype=np.float32)
    wav = (360 + 359 * rng.random((4, 513), dtype=np.float32)).astype(np.float32) if spectral else None
    calls = [
        [dict(turbidity=2.2, albedo=0.15, sun_direction=sw.sun_direction(12.0)),
         dict(turbidity=7.5, albedo=0.8, sun_direction=sw.sun_direction(64.0))],
        [dict(turbidity=9.0, albedo=0.0, sun_direction=sw.sun_direction(3.0)),
         dict(turbidity=1.3, albedo=1.0, sun_direction=sw.sun_direction(85.0))],
    ]
    seen = []
    for props in calls:                                         # two calls in a row on the same shared scratch
        assert _call(L, "ssk_update_batch_noalloc", a, props) == 0, L.ssk_last_error().decode()
        assert _call(L, "ssk_update_batch", b, props) == 0, L.ssk_last_error().decode()
        for x, y in zip(a, b):
            sx, sy = x.state(), y.state()
            assert bytes(sx) == bytes(sy), "ssk_state differs between the no-alloc and the allocating update"
            assert common.bit_equal(np.float32(sx.sky_sampling_w), np.float32(sy.sky_sampling_w))
            assert common.bit_equal(x.eval(-wo, wav), y.eval(-wo, wav))
            dx, px, wx = x.sample_direction(u, None, wav); dy, py, wy = y.sample_direction(u, None, wav)
            assert common.bit_equal(dx, dy) and common.bit_equal(px, py) and common.bit_equal(wx, wy)
        seen.append([float(x.state().sky_sampling_w) for x in a])
    assert seen[0] != seen[1], "the second call must have produced new state, not the first call's"
    # a single emitter (n below the maximum) after a full batch: stale partials of slot 1 must not matter
    one = [dict(turbidity=4.4, albedo=0.5, sun_direction=sw.sun_direction(41.0))]
    assert _call(L, "ssk_update_batch_noalloc", a[:1], one) == 0, L.ssk_last_error().decode()
    assert _call(L, "ssk_update_batch", b[:1], one) == 0
    assert bytes(a[0].state()) == bytes(b[0].state())
    assert bytes(a[1].state()) == bytes(b[1].state())          # untouched by the one-emitter call


def test_noalloc_update_rejects_what_it_cannot_hold(ssk):
    from sunsky_b200 import _capi
    L = _capi.lib()
    es = [ssk.SunskyEmitter(turbidity=3.0, albedo=0.3) for _ in range(3)]
    before = [bytes(e.state()) for e in es]
    assert _call(L, "ssk_update_batch_noalloc", es, [dict(turbidity=5.0)] * 3) != 0
    msg = L.ssk_last_error().decode()
    assert "SSK_NOALLOC_MAX_EMITTERS" in msg and "3 emitters" in msg
    # an invalid property: error, nothing modified
    assert _call(L, "ssk_update_batch_noalloc", es[:2], [dict(turbidity=5.0), dict(turbidity=11.0)]) != 0
    assert [bytes(e.state()) for e in es] == before
    # n = 0 is a no-op
    assert L.ssk_update_batch_noalloc(None, None, 0) == 0
