import sys, time
sys.path.insert(0, 'mitsuba3-sunsky_b200'); sys.path.insert(0, 'tests')
import sunsky_b200 as ssk, common, torch
for sp in (False, True):
    e = ssk.SunskyEmitter(spectral=sp, turbidity=3.0, albedo=0.3, sun_direction=common.sun_dir_from_elevation(30.0))
    for k in range(4): e.update(turbidity=3.0 + 0.1 * k)
    t0 = time.perf_counter()
    for k in range(100): e.update(turbidity=3.0 + 0.01 * k)
    print('spectral' if sp else 'rgb', 'update ms', (time.perf_counter() - t0) * 10, file=sys.stderr)
