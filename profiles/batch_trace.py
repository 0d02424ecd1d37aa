import sys, time
sys.path.insert(0, 'mitsuba3-sunsky_b200'); sys.path.insert(0, 'tests')
import sunsky_b200 as ssk, sunsky_b200.sweep as sw, torch
for sp in (False, True):
    for n in (8, 16, 32):
        es = [ssk.SunskyEmitter(spectral=sp, turbidity=3.0, albedo=0.3, sun_direction=sw.sun_direction(30.0)) for _ in range(n)]
        props = [dict(turbidity=2.0 + 0.1 * k, albedo=0.3, sun_direction=sw.sun_direction(10.0 + k)) for k in range(n)]
        ssk.update_batch(es, props)
        t0 = time.perf_counter()
        for _ in range(10): ssk.update_batch(es, props)
        print('spectral' if sp else 'rgb', n, 'ms per emitter', (time.perf_counter() - t0) / 10 / n * 1e3, file=sys.stderr)
        for e in es: e.close()
