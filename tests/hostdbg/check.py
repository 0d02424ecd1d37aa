"""DEBUG AID (container only): host execution of the per-lane model vs the golden vectors."""
import sys, numpy as np, ctypes as C, os
HERE = os.path.dirname(os.path.abspath(__file__)); REPO = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, os.path.join(REPO, 'tests')); sys.path.insert(0, os.path.join(REPO, 'mitsuba3-sunsky_b200'))
import common
from sunsky_b200 import _capi as A
H = C.CDLL(os.path.join(HERE, 'libhostdbg.so')); H.hostdbg_create.restype = C.c_void_p
FP = A.FP
fp = lambda a: a.ctypes.data_as(FP)
L = A.lib()
m = np.load(os.path.join(REPO, 'tests/golden/drjit_math.npz'))
ops = dict(sin=0, cos=1, asin=2, acos=3, atan2=4, exp=5, exp2=6, log=7, log2=8, cbrt=9, erf=10, erfinv=11, pow15=12, tan=13, sqrt=14)
for op, code in ops.items():
    if op == 'atan2':
        a = np.ascontiguousarray(m['atan2_y']); b = np.ascontiguousarray(m['atan2_x'])
    else:
        a = np.ascontiguousarray(m[op + '_x']); b = a
    o = np.empty_like(a)
    H.hostdbg_math(code, C.c_size_t(a.size), fp(a), fp(b), fp(o))
    ref = m[op]
    same = (o.view(np.uint32) == ref.view(np.uint32)) | (np.isnan(o) & np.isnan(ref))
    bad = np.nonzero(~same)[0]
    print(f"{op:8s} mismatches {bad.size}/{a.size}", [(a[i], o[i], ref[i]) for i in bad[:3]])
only = sys.argv[1:] 
for name, (spectral, kw) in common.CONFIGS.items():
    if only and name not in only: continue
    g = np.load(os.path.join(REPO, f'tests/golden/{name}.npz'))
    p = A.Params(); L.ssk_params_default(C.byref(p))
    p.turbidity = kw['turbidity']
    alb = np.atleast_1d(np.asarray(kw['albedo'], dtype=np.float32)); p.n_albedo = alb.size
    for i, v in enumerate(alb): p.albedo[i] = v
    if 'sun_direction' in kw:
        p.use_record = 0
        for i in range(3): p.sun_direction[i] = g['state_sun_dir_world'][i]
        p.flags = 1
    else:
        p.hour = kw['hour']
    p.sun_scale = kw.get('sun_scale', 1.0); p.arith = int(os.environ.get('SSK_ARITH', '0'))
    if 'sun_aperture' in kw: p.sun_aperture_deg = kw['sun_aperture']
    h = H.hostdbg_create(C.byref(p), A.DATASET_DIR.encode(), int(spectral))
    wo = np.ascontiguousarray(g['wo']); n = wo.shape[1]; wi = np.ascontiguousarray(-wo)
    wav = np.ascontiguousarray(g['wav']) if spectral else None
    Cn = 4 if spectral else 3
    out = np.empty((Cn, n), np.float32)
    H.hostdbg_eval(C.c_void_p(h), C.c_size_t(n), fp(wi), fp(wav) if spectral else None, fp(out))
    pdf = np.empty(n, np.float32); H.hostdbg_pdf(C.c_void_p(h), C.c_size_t(n), fp(wo), fp(pdf))
    u = np.ascontiguousarray(g['u']); d = np.empty((3, n), np.float32); sp = np.empty(n, np.float32); sw = np.empty((Cn, n), np.float32)
    H.hostdbg_sample(C.c_void_p(h), C.c_size_t(n), fp(u), fp(wav) if spectral else None, fp(d), fp(sp), fp(sw))
    def mm(a, b):
        same = (a.view(np.uint32) == b.view(np.uint32)) | (np.isnan(a) & np.isnan(b)) | ((a == 0) & (b == 0))
        return int((~same).sum())
    print(f"{name:18s} eval mism {mm(out, g['eval'])}  pdf mism {mm(pdf, g['pdf'])}  d mism {mm(d, g['s_d'])}  spdf mism {mm(sp, g['s_pdf'])} weight mism {mm(sw, g['s_weight'])}   maxrel eval {common.max_rel_err(out, g['eval'])} pdf {common.max_rel_err(pdf, g['pdf'])}")
    wl = np.ascontiguousarray(g['wl_sample']); s2 = np.ascontiguousarray(g['s2'])
    if spectral:
        ww = np.empty((4, n), np.float32); wq = np.empty((4, n), np.float32)
        H.hostdbg_sample_wavelengths(C.c_void_p(h), C.c_size_t(n), fp(wi), fp(wl), fp(ww), fp(wq))
        print(f"{'':18s} sample_wavelengths: wav mism {mm(ww, g['sw_wav'])} weight mism {mm(wq, g['sw_weight'])} maxrel {common.max_rel_err(wq, g['sw_weight'])}")
    H.hostdbg_set_bsphere(C.c_void_p(h), fp(np.ascontiguousarray(g['state_bsphere'])))
    ro = np.empty((3, n), np.float32); rd = np.empty((3, n), np.float32); rw = np.zeros((4, n), np.float32); rq = np.empty((Cn, n), np.float32)
    H.hostdbg_sample_ray(C.c_void_p(h), C.c_size_t(n), fp(wl), fp(s2), fp(u), fp(ro), fp(rd), fp(rw), fp(rq))
    print(f"{'':18s} sample_ray: o mism {mm(ro, g['ray_o'])} d mism {mm(rd, g['ray_d'])} wav mism {mm(rw, g['ray_wav'])} weight mism {mm(rq, g['ray_weight'])} maxrel {common.max_rel_err(rq, g['ray_weight'])}")
    if only:
        np.savez('/tmp/hostdbg_ray_%s.npz' % name, ro=ro, rd=rd, rw=rw, rq=rq)
    if only:
        np.savez('/tmp/hostdbg_%s.npz' % name, out=out, pdf=pdf, d=d, sp=sp, sw=sw)
