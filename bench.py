#!/usr/bin/env python
"""Benchmark of the sunsky hot path (see DESIGN.md "Measurement").

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]

One JSON line on stdout. A "step" is one pass of SunskyEmitter.eval (spectral, sun + sky) over the
BASELINE configs[1] batch: 2^24 upper-hemisphere directions with 4 wavelengths each, drawn from PCG32 streams as
SURVEY.md 8(d) cfg 2 lays them out, resident in HBM (470 MB of inputs per step, larger than the 126 MB L2, so no
flush is needed between steps). `value` = Gevals/s over all ranks (device time, CUDA events, max over ranks);
`e2e` = the same through ssk_eval_host with pinned host buffers (H2D + kernel + D2H inside the timed region), next
to the bare-copy ceiling of the same bytes; `sweeps` = BASELINE configs[3] / [4] (with the frames gathered on rank 0
when N > 1); `extra` carries the other kernels of the path (RGB eval, pdf_direction, fused sample_direction on the
configs[2] batch) and BASELINE configs[0]. `--impl reference` times the reference's own CPU code (oracle/_ref) on
the same batch with all host threads."""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

REPO = os.path.dirname(os.path.abspath(__file__))
for p in (os.path.join(REPO, "mitsuba3-sunsky_b200"), os.path.join(REPO, "oracle"), os.path.join(REPO, "tests")):
    if p not in sys.path:
        sys.path.insert(0, p)

import numpy as np  # noqa: E402

N_EVAL = 1 << 24            # BASELINE configs[1]: 16M directions
N_SAMPLE = 1 << 26          # BASELINE configs[2]: 64M samples
FLOPS = dict(eval_spectral=1013, eval_rgb=426, pdf=761, sample_rgb=1776)
BYTES = dict(eval_spectral=44, eval_rgb=24, pdf=16, sample_rgb=36)
METRIC = "sky+sun radiance Gevals/s (spectral eval, 16M directions)"
WORKLOAD = "configs[1]: spectral eval, 2^24 upper-hemisphere directions x 4 wavelengths per GPU, t=3 a=0.3 eta=30deg, sun+sky"
# identical in both arms (the driver compares them)
CONFIG = {"workload": WORKLOAD, "l2": "inputs (470 MB per step) exceed the 126 MB L2; no flush needed between steps"}
# Constants read off the committed ncu --set full captures of eval_spec_kernel on 2^24 rays (profiles/NCU_SOURCE):
# dram__bytes_read.sum + dram__bytes_write.sum per launch (algorithmic: 44 B x 2^24 = 738 MB; part of the output is
# still in L2 when the kernel ends), issue-slot utilisation, and executed fp32 flop per ray (FFMA = 2, FADD / FMUL = 1,
# packed f32x2 ops counted per element, MUFU listed apart). They are NOT measured in this run; the key says so.
NCU_SOURCE = "profiles/r2_summary.md (r2_ncu_full_default_*_eval_spec.txt, r2_sassmix_default_*_eval_spec.txt)"
NCU = {"fast": dict(traffic=702.0e6, issue_frac=0.7185, flop_per_ray=361, mufu_per_ray=28),
       "exact": dict(traffic=699.7e6, issue_frac=0.6678, flop_per_ray=1163, mufu_per_ray=12)}


def peaks():
    p = dict(hbm_gbs=6650.0, sm_max_mhz=1965.0, source="fallback")
    try:
        with open(os.path.join(REPO, "MEASURED_PEAKS.json")) as f:
            m = json.load(f)
        p.update(hbm_gbs=float(m["hbm_gbs"]), sm_max_mhz=float(m["sm_max_mhz"]), source="measured")
    except Exception:
        pass
    return p


class ClockSampler:
    """SM clock and throttle reasons of one GPU, polled through NVML from a thread while the timed regions
    run (nvidia-smi's own loop mode starts too slowly for a sub-second region). Samples carry a host
    timestamp so the ones inside a timed region can be told apart from the rest of the run."""
    REASONS = (("hw_slowdown", 0x8), ("hw_thermal_slowdown", 0x40), ("sw_thermal_slowdown", 0x20), ("sw_power_cap", 0x4))

    def __init__(self, uuid, period_s=0.01):
        self.uuid = uuid; self.period = period_s; self.rows = []; self.h = None; self.stop_flag = False
        self.max_mhz = None; self.t = None; self.nv = None

    def start(self):
        try:
            import pynvml as nv
            nv.nvmlInit()
            try:
                self.h = nv.nvmlDeviceGetHandleByUUID(self.uuid.encode() if isinstance(self.uuid, str) else self.uuid)
            except Exception:
                self.h = nv.nvmlDeviceGetHandleByUUID(self.uuid)
            self.nv = nv
            self.max_mhz = float(nv.nvmlDeviceGetMaxClockInfo(self.h, nv.NVML_CLOCK_SM))
            self.t = threading.Thread(target=self._poll, daemon=True); self.t.start()
        except Exception as ex:
            self.err = str(ex); self.h = None

    def _poll(self):
        nv = self.nv
        while not self.stop_flag:
            try:
                mhz = nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM)
                try:
                    rs = nv.nvmlDeviceGetCurrentClocksEventReasons(self.h)
                except Exception:
                    rs = nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.h)
                try:
                    pw = nv.nvmlDeviceGetPowerUsage(self.h) / 1e3
                except Exception:
                    pw = None
                self.rows.append((time.perf_counter(), float(mhz), int(rs), pw))
            except Exception:
                pass
            time.sleep(self.period)

    def stop(self):
        self.stop_flag = True
        if self.t:
            self.t.join(timeout=1)

    def summary(self, windows):
        """Median SM clock over the samples that fall inside any (t0, t1) window; falls back to the whole
        run when the windows are too short to hold a sample."""
        inside = [r for r in self.rows if any(a <= r[0] <= b for a, b in windows)]
        where = "timed regions"
        if not inside:
            inside = self.rows; where = "whole run (timed regions shorter than the polling period)"
        reasons = set()
        for r in inside:
            for name, bit in self.REASONS:
                if r[2] & bit:
                    reasons.add(name)
        pw = [r[3] for r in inside if r[3] is not None]
        return dict(sm_mhz=float(np.median([r[1] for r in inside])) if inside else None, sm_max_mhz=self.max_mhz,
                    reasons=sorted(reasons), samples=len(inside), window=where,
                    power_w_max=max(pw) if pw else None)


def emitter_kwargs():
    import common
    return dict(turbidity=3.0, albedo=0.3, sun_direction=common.sun_dir_from_elevation(30.0))


def configs1_inputs(n, device, rank=0):
    from sunsky_b200 import pcg32
    return pcg32.configs1_inputs(n, device, rank)


def cpu_rate(r, wi, wav, threads, n, repeat=1):
    """Gevals/s and seconds of the reference emitter `r` over the first n lanes (numpy planes), best of `repeat`."""
    wi_n = np.ascontiguousarray(wi[:, :n]); wav_n = None if wav is None else np.ascontiguousarray(wav[:, :n])
    best = 1e30
    for _ in range(repeat):
        t0 = time.perf_counter()
        r.eval(wi_n, wav_n, threads=threads)
        best = min(best, time.perf_counter() - t0)
    return n / best / 1e9, best


def cpu_flavours(wi, wav, threads, budget_s=24.0):
    """The reference's CPU flavours on (a bounded sample of) the configs[1] batch: scalar (= scalar_spectral) and
    packets (its query members on AVX2 / AVX-512 packets, the stand-in for llvm_ad_spectral's vectorised kernels),
    one thread and all threads. Sample sizes are chosen from a calibration pass so the whole call stays near
    `budget_s` seconds of CPU work."""
    import ref_oracle
    isa = ref_oracle.best_isa()
    kw = emitter_kwargs()
    out = {"isa": isa, "packet_width": int(ref_oracle.lib(isa).ssk_ref_packet_width()), "threads": threads}
    n_max = wi.shape[1]
    for name, flavour in (("scalar", 0), ("packet", 2)):
        r = ref_oracle.RefEmitter(spectral=True, flavour=flavour, isa=isa, **kw)
        r.eval(wi[:, :4096].copy(), wav[:, :4096].copy(), threads=1)
        cal, _ = cpu_rate(r, wi, wav, threads, min(n_max, 1 << 18))
        n_all = int(min(n_max, max(1 << 18, 2 ** int(np.log2(cal * 1e9 * budget_s / 6)))))
        n_one = int(min(n_max, max(1 << 16, n_all // max(threads, 1))))
        # at least ~3 s of repeated passes: the first passes of a freshly started thread pool run 20-30 % slow
        _, t_first = cpu_rate(r, wi, wav, threads, n_all)
        v_all, t_all = cpu_rate(r, wi, wav, threads, n_all, repeat=int(min(40, max(3, np.ceil(3.0 / max(t_first, 1e-3))))))
        v_one, t_one = cpu_rate(r, wi, wav, 1, n_one)
        out[name] = {"value": v_all, "single_thread_value": v_one, "lanes": n_all, "seconds": t_all,
                     "single_thread_lanes": n_one}
    return out


def run_reference(args):
    """--impl reference: the reference's CPU implementation of the path on the host cores (rank 0 only). The timed
    flavour is the fastest one the reference's code has on this box: its eval() on dr::Packet<float, 16 | 8> (what
    llvm_ad_spectral executes per SIMD lane group, minus the JIT) on every host thread. Every step is the FULL
    configs[1] batch (2^24 directions) unless a calibration pass shows that warmup + steps would take more than
    about three minutes; then the step shrinks to the largest power of two that fits, and `ms_per_step` is the
    measured time of THAT step, never an extrapolation."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return 0
    import ref_oracle
    if not ref_oracle.available():
        print(json.dumps({"impl": "reference", "unavailable": "oracle/_ref/libsunsky_ref.so was not built"}))
        return 0
    import torch
    threads = min(os.cpu_count() or 1, 64)
    torch.set_num_threads(threads)
    wi_t, _, wav_t = configs1_inputs(N_EVAL, "cpu", 0)
    wi, wav = wi_t.numpy(), wav_t.numpy()
    isa = ref_oracle.best_isa()
    width = int(ref_oracle.lib(isa).ssk_ref_packet_width())
    r = ref_oracle.RefEmitter(spectral=True, flavour=2, isa=isa, **emitter_kwargs())
    cal, _ = cpu_rate(r, wi, wav, threads, 1 << 20)
    n_step = N_EVAL
    while n_step > (1 << 18) and (args.warmup + args.steps) * n_step / (cal * 1e9) > 170.0:
        n_step //= 2
    times = []
    for i in range(args.warmup + args.steps):
        _, dt = cpu_rate(r, wi, wav, threads, n_step)
        if i >= args.warmup:
            times.append(dt)
    total_t = sum(times)
    value = n_step * len(times) / total_t / 1e9
    rs = ref_oracle.RefEmitter(spectral=True, flavour=0, isa=isa, **emitter_kwargs())
    v_scalar, _ = cpu_rate(rs, wi, wav, threads, 1 << 20)
    sample = (f"{n_step} of the {N_EVAL} directions per step" + (" (the full batch)" if n_step == N_EVAL else " (calibrated to the time budget)") +
              f"; the reference's eval() on dr::Packet<float, {width}> ({isa}) over {threads} host threads; its scalar_spectral "
              f"flavour on the same threads: {v_scalar * 1e3:.2f} Mevals/s")
    print(json.dumps({
        "impl": "reference", "metric": METRIC, "value": value, "unit": "Gevals/s", "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * total_t / len(times),
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic (PCG32 streams, SURVEY.md 8d cfg 2)",
        "config": CONFIG,
        "cpu_baseline": {"value": value, "unit": "Gevals/s", "cores": threads, "kind": "reference", "sample": sample,
                         "isa": isa, "packet_width": width, "scalar_value": v_scalar, "lanes_per_step": n_step},
        "e2e": {"value": value, "unit": "Gevals/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }))
    return 0


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=300)
    ap.add_argument("--warmup", type=int, default=10)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--no-extra", action="store_true", help="skip the secondary kernels and the CPU baseline")
    ap.add_argument("--sweeps-only", action="store_true", help="tuning runs: headline + env-map sweeps, no secondary kernels")
    ap.add_argument("--arith", default="fast", choices=["fast", "exact"],
                    help="arithmetic mode of the headline kernel: fast (within 1e-5 of the reference, default) or exact "
                         "(bit-identical to the reference)")
    args = ap.parse_args()
    if args.impl == "reference":
        return run_reference(args)

    import torch
    import torch.distributed as dist
    import sunsky_b200 as ssk

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device: the sunsky kernels have no CPU fallback")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def max_over_ranks(ms):
        if world > 1:
            t = torch.tensor([ms], device=dev, dtype=torch.float64)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            return float(t.item())
        return ms

    warm = max(args.warmup, 3)

    windows = []                                    # host-time (t0, t1) of every timed region, for the clock record

    def timed(fn, steps, warmup):
        for _ in range(warmup):
            fn()
        barrier()
        ev0 = torch.cuda.Event(enable_timing=True); ev1 = torch.cuda.Event(enable_timing=True)
        l0 = ssk.launch_count()
        t0 = time.perf_counter()
        ev0.record()
        for _ in range(steps):
            fn()
        ev1.record()
        barrier()
        windows.append((t0, time.perf_counter()))
        return max_over_ranks(ev0.elapsed_time(ev1)), ssk.launch_count() - l0

    kw = emitter_kwargs()
    pk = peaks()
    sm_count = torch.cuda.get_device_properties(dev).multi_processor_count
    fp32_peak = sm_count * 128 * 2 * pk["sm_max_mhz"] * 1e6 / 1e12          # TFLOP/s at max SM clock
    import ctypes as C
    from sunsky_b200 import _capi
    import common
    L = _capi.lib()
    stream = C.c_void_p(torch.cuda.current_stream().cuda_stream)

    def planes(t):
        arr = (_capi.FP * t.shape[0])()
        for i in range(t.shape[0]):
            arr[i] = C.cast(t[i].data_ptr(), _capi.FP)
        return arr

    def fp(t):
        return C.cast(t.data_ptr(), _capi.FP)

    # ---- inputs, resident in HBM ----------------------------------------------------------------------
    wi, wo, wav = configs1_inputs(N_EVAL, dev, rank)
    out = torch.empty((4, N_EVAL), device=dev)
    wi_pp, wo_pp, wav_pp, out_pp = planes(wi), planes(wo), planes(wav), planes(out)

    def rate(ms_total, steps, n, key):
        r = world * n / (ms_total / steps * 1e-3) / 1e9
        return {"value": r, "unit": "G/s", "ms": ms_total / steps,
                # the reference's op count at this rate, over the fp32 peak: > 1 where FAST mode or
                # hoisted constants skip work the reference does -- a reference-equivalent rate, not pipe utilisation
                "fp32_frac_algorithmic": r / world * 1e9 * FLOPS[key] / 1e12 / fp32_peak,
                "hbm_gbs": r / world * BYTES[key], "hbm_frac": r / world * BYTES[key] / pk["hbm_gbs"]}

    # ---- headline: spectral eval (configs[1]) in the arithmetic mode under test --------------------------
    em_spec = ssk.SunskyEmitter(spectral=True, device=local, arith=args.arith, **kw)

    def step_eval():
        rc = L.ssk_eval_device(em_spec._ctx, wi_pp, wav_pp, N_EVAL, out_pp, stream)
        assert rc == 0, L.ssk_last_error()

    sampler = ClockSampler("GPU-" + str(torch.cuda.get_device_properties(dev).uuid)); sampler.start()
    ms, launches = timed(step_eval, args.steps, warm)
    headline_window = list(windows)
    ms_per_step = ms / args.steps
    value = world * N_EVAL / (ms_per_step * 1e-3) / 1e9
    per_gpu = value / world
    achieved_tf = per_gpu * 1e9 * FLOPS["eval_spectral"] / 1e12
    ncu = NCU[args.arith]
    exec_tf = None if not ncu["flop_per_ray"] else per_gpu * 1e9 * ncu["flop_per_ray"] / 1e12
    roofline = {"bound": "fp32", "achieved": achieved_tf, "peak": fp32_peak, "unit": "TFLOP/s",
                "frac": achieved_tf / fp32_peak, "traffic": ncu["traffic"],
                # what the hardware executed, as opposed to what the reference's op sequence would cost
                "issue_frac": ncu["issue_frac"],
                "executed": {"flop_per_ray": ncu["flop_per_ray"], "mufu_per_ray": ncu["mufu_per_ray"],
                             "achieved": exec_tf, "frac": None if exec_tf is None else exec_tf / fp32_peak, "unit": "TFLOP/s",
                             "source": f"constants from the ncu --set full capture in {NCU_SOURCE}, not measured in this run"},
                "peak_source": f"{sm_count} SMs x 128 lanes x 2 flop x {pk['sm_max_mhz']:.0f} MHz ({pk['source']} max SM clock); "
                               "not a tensor-core path",
                "kernel": f"eval_spec_kernel<{args.arith}>", "flops_per_eval": FLOPS["eval_spectral"],
                "flops_note": "algorithmic flops of the reference op sequence (SURVEY.md 8d), not executed instructions",
                "traffic_note": f"dram read+write bytes per launch, constant from the ncu capture in {NCU_SOURCE} (algorithmic: 738 MB)",
                "hbm": {"achieved": per_gpu * BYTES["eval_spectral"], "peak": pk["hbm_gbs"], "unit": "GB/s",
                        "frac": per_gpu * BYTES["eval_spectral"] / pk["hbm_gbs"], "bytes_per_eval": BYTES["eval_spectral"]}}

    # ---- e2e: same workload through the host entry point with pinned buffers ------------------------
    n_e2e = N_EVAL
    h_in = torch.empty((7, n_e2e), pin_memory=True); h_in[:3].copy_(wi); h_in[3:].copy_(wav)     # si.wi | si.wavelengths
    h_out = torch.empty((4, n_e2e), pin_memory=True)
    hwi_pp, hwav_pp, hout_pp = planes(h_in[:3]), planes(h_in[3:]), planes(h_out)

    def step_e2e():
        rc = L.ssk_eval_host(em_spec._ctx, hwi_pp, hwav_pp, n_e2e, hout_pp)
        assert rc == 0, L.ssk_last_error()

    e2e_steps = max(3, min(args.steps, 20))
    for _ in range(2):
        step_e2e()
    barrier()
    t0 = time.perf_counter()
    l0 = ssk.launch_count()
    for _ in range(e2e_steps):
        step_e2e()
    barrier()
    windows.append((t0, time.perf_counter()))
    e2e_s = (windows[-1][1] - t0) / e2e_steps
    e2e_ms = max_over_ranks(e2e_s * 1e3)
    # the bare transfers of the same bytes, all ranks at once: what the host <-> device links of this box allow
    ceil_s, ceil_mean_s = C.c_double(), C.c_double()
    barrier()
    rc = L.ssk_host_copy_ceiling(local, 7 * 4 * n_e2e, 4 * 4 * n_e2e, 7 * 4 * (1 << 19), 8, C.byref(ceil_s), C.byref(ceil_mean_s))
    assert rc == 0, L.ssk_last_error()
    barrier()
    ceil_best_ms = max_over_ranks(ceil_s.value * 1e3)
    ceil_ms = max_over_ranks(ceil_mean_s.value * 1e3)          # every rank copies at the same time: the contended ceiling
    e2e_bytes = (7 + 4) * 4 * n_e2e
    e2e = {"value": world * n_e2e / (e2e_ms * 1e-3) / 1e9, "unit": "Gevals/s",
           "h2d_bytes_per_step": 7 * 4 * n_e2e, "d2h_bytes_per_step": 4 * 4 * n_e2e, "ms_per_step": e2e_ms,
           "launches_per_step": (ssk.launch_count() - l0) // e2e_steps,
           "pcie_gbs": e2e_bytes / (e2e_ms * 1e-3) / 1e9,
           "pcie_peak_gbs": e2e_bytes / (ceil_ms * 1e-3) / 1e9, "copy_ceiling_ms": ceil_ms, "frac": ceil_ms / e2e_ms,
           "copy_ceiling_best_ms": ceil_best_ms,
           "note": "ssk_eval_host: pinned host planes, 512 Ki-ray chunks over 3 streams, one strided H2D and one strided D2H copy per "
                   "chunk, kernel in between, all inside the timed region. pcie_peak_gbs: the same bytes as bare pinned copies "
                   "(H2D on one stream, D2H on another, every rank at once; ssk_host_copy_ceiling), per GPU: mean of 8 passes, max "
                   "over ranks (copy_ceiling_best_ms: the best pass, which at N > 1 is a pass some other rank had already "
                   "left); frac = ceiling time / e2e time"}
    del h_in, h_out

    # ---- the other kernels of the path, both arithmetic modes ----------------------------------------------
    extra = {}
    if not args.no_extra and not args.sweeps_only:
        out3 = torch.empty((3, N_EVAL), device=dev); o3_pp = planes(out3)
        pdf = torch.empty(N_EVAL, device=dev)
        from sunsky_b200 import pcg32
        u = pcg32.point2_pairs(N_SAMPLE, stream_offset=(1 << 40) + rank * 2 * N_SAMPLE, device=dev)   # sampling_test.py:30-32
        d = torch.empty((3, N_SAMPLE), device=dev); spdf = torch.empty(N_SAMPLE, device=dev)
        wgt = torch.empty((3, N_SAMPLE), device=dev)
        u_pp, d_pp, w_pp = planes(u), planes(d), planes(wgt)
        s_steps = max(3, args.steps // 10)
        for mode in ("fast", "exact"):
            ex = {}
            es = ssk.SunskyEmitter(spectral=True, device=local, arith=mode, **kw)
            er = ssk.SunskyEmitter(spectral=False, device=local, arith=mode, **kw)
            # configs[2] = sunsky-testing/sampling_test.py: t=6, a=0.5, eta=50.2 deg, sky only; and its sun+sky twin
            kw3 = dict(turbidity=6.0, albedo=0.5, sun_direction=common.sun_dir_from_elevation(50.2))
            e_sky = ssk.SunskyEmitter(spectral=False, device=local, arith=mode, sun_scale=0.0, **kw3)
            e_sun = ssk.SunskyEmitter(spectral=False, device=local, arith=mode, **kw3)

            def step_spec():
                assert L.ssk_eval_device(es._ctx, wi_pp, wav_pp, N_EVAL, out_pp, stream) == 0

            def step_rgb():
                assert L.ssk_eval_device(er._ctx, wi_pp, None, N_EVAL, o3_pp, stream) == 0

            def step_pdf():
                assert L.ssk_pdf_direction_device(er._ctx, wo_pp, N_EVAL, fp(pdf), stream) == 0

            def step_sample(e):
                def f():
                    assert L.ssk_sample_direction_device(e._ctx, u_pp, None, None, N_SAMPLE, d_pp, fp(spdf), w_pp,
                                                         None, None, stream) == 0
                return f
            if mode != args.arith:
                ms_x, _ = timed(step_spec, args.steps, warm)
                ex["eval_spectral_Gevals_s"] = rate(ms_x, args.steps, N_EVAL, "eval_spectral")
            ms_r, _ = timed(step_rgb, args.steps, warm)
            ex["eval_rgb_Gevals_s"] = rate(ms_r, args.steps, N_EVAL, "eval_rgb")
            ms_p, _ = timed(step_pdf, args.steps, warm)
            ex["pdf_direction_Gpdf_s"] = rate(ms_p, args.steps, N_EVAL, "pdf")
            ms_s, _ = timed(step_sample(e_sky), s_steps, 3)
            ex["sample_direction_sky_only_Gsamples_s"] = rate(ms_s, s_steps, N_SAMPLE, "sample_rgb")
            ex["sample_direction_lower_hemisphere"] = int((d[2] < 0).sum().item())      # sampling_test.py:36-44
            ms_s, _ = timed(step_sample(e_sun), s_steps, 3)
            ex["sample_direction_sun_and_sky_Gsamples_s"] = rate(ms_s, s_steps, N_SAMPLE, "sample_rgb")
            extra[mode] = ex
            for e in (es, er, e_sky, e_sun):
                e.close()
        del out3, pdf, u, d, spdf, wgt
        torch.cuda.empty_cache()

    # ---- configs[3] / configs[4]: env-map sweeps through ssk_sweep_render (C++ loop: pooled contexts re-parametrised
    #      in batches, setup quadrature overlapping the queued maps). Each rank renders its share of the frames; with
    #      N > 1 the frames are also collected on rank 0, (a) by the kernels themselves storing into rank 0's
    #      IPC-mapped buffer over NVLink and (b) by the library's NCCL send/recv gather on a second stream.
    sweeps = None
    if not args.no_extra:
        from sunsky_b200 import multi, sharding, sweep as sweep_mod
        NVLINK_INGEST_GBS = 900.0        # NVLink 5 per direction per GPU (nominal): floor of all-frames-to-one-GPU
        comm_box = []                    # one communicator for all NCCL sweeps: peer connections are set up on first use

        def sweep_bench(config, width, height, chunk, gather):
            spectral = config == "animation"
            mode_ = multi.SHARD_ROUND_ROBIN if spectral else multi.SHARD_BLOCKED
            allp = sharding.sweep_parameters(config)
            n_frames = len(allp)
            mine = multi.frames_for_rank(n_frames, rank, world, mode_)
            props = [dict(turbidity=allp[f]["turbidity"], albedo=allp[f]["albedo"],
                          sun_direction=sweep_mod.sun_direction(allp[f]["eta_deg"])) for f in mine]
            fparams = multi.frame_params(props, spectral, args.arith)
            Cn = 4 if spectral else 3
            frame_floats = Cn * width * height
            wl = (440.0, 520.0, 600.0, 680.0)
            peer = comm = full = staging = buf = None
            if gather == "peer":
                peer = multi.PeerFrames(n_frames * frame_floats, local, rank, 0, dist if world > 1 else None)
                allptr = peer.frame_pointers(frame_floats, n_frames)
                ptrs = (_capi.FP * max(len(mine), 1))(*[allptr[f] for f in mine])
            elif gather == "nccl":
                if not comm_box:
                    comm_box.append(multi.Comm(rank, world, local, dist))
                comm = comm_box[0]
                staging = torch.empty((2 * chunk, Cn, height, width), device=dev)
                full = torch.empty((n_frames, Cn, height, width), device=dev) if rank == 0 else None
            else:                                    # frames stay where they are rendered (a ring of 2 chunks is kept)
                buf = torch.empty((2 * chunk, Cn, height, width), device=dev)
                ptrs = (_capi.FP * max(len(mine), 1))(*[C.cast(buf[k % (2 * chunk)].data_ptr(), _capi.FP) for k in range(len(mine))])
            l0 = ssk.launch_count()
            barrier()
            t0 = time.perf_counter()
            if gather == "nccl":
                wl_np = np.ascontiguousarray(np.broadcast_to(np.asarray(wl, np.float32), (max(len(mine), 1), 4)))
                rc = L.ssk_sweep_render_gather_nccl(comm.h, fparams, n_frames, mode_, _capi.DATASET_DIR.encode(),
                                                    _capi.MODE_SPECTRAL if spectral else _capi.MODE_RGB, width, height,
                                                    wl_np.ctypes.data_as(_capi.FP) if spectral else None, fp(staging),
                                                    fp(full) if rank == 0 else None, 0, chunk, stream)
                assert rc == 0, L.ssk_multi_last_error()
            else:
                multi.sweep_render(fparams, spectral, width, height, local, ptrs, wavelengths4=wl, chunk=chunk)
            torch.cuda.synchronize()
            barrier()
            windows.append((t0, time.perf_counter()))
            sec = max_over_ranks(windows[-1][1] - t0)
            res = {"frames": n_frames, "frames_per_gpu": len(mine), "resolution": [width, height], "seconds": sec,
                   "frames_per_s": n_frames / sec, "Gpixels_s": n_frames * width * height / sec / 1e9,
                   "gpu_launches": ssk.launch_count() - l0, "gather": gather or "none",
                   "out_GB_total": n_frames * frame_floats * 4 / 1e9}
            if gather:
                remote = (n_frames - len(multi.frames_for_rank(n_frames, 0, world, mode_))) * frame_floats * 4
                floor = remote / (NVLINK_INGEST_GBS * 1e9)
                res.update(floor_seconds=floor, seconds_over_floor=sec / floor if floor > 0 else None,
                           ingest_GB_s=remote / sec / 1e9,
                           floor_note=f"bytes that must enter rank 0 ({remote / 1e9:.1f} GB) / {NVLINK_INGEST_GBS:.0f} GB/s NVLink ingest of one GPU")
                if rank == 0:                        # the gathered frames are really there: spot-check one remote frame
                    got = peer.tensor((n_frames, Cn, height, width)) if gather == "peer" else full
                    res["checksum_last_frame"] = float(got[n_frames - 1].double().sum().item())
            if peer is not None:
                peer.close()
            del full, staging, buf
            torch.cuda.empty_cache()
            return res

        sw = {}
        sweep_bench("time_of_day", 1024, 512, 8, None)                      # warm-up: contexts, scratch
        sw["time_of_day_256x4096x2048_rgb"] = sweep_bench("time_of_day", 4096, 2048, 8, None)
        sw["animation_128x8192x4096_spectral"] = sweep_bench("animation", 8192, 4096, 2, None)
        if world > 1:
            for how in ("peer", "nccl"):
                sweep_bench("time_of_day", 1024, 512, 8, how)
                sw[f"time_of_day_256x4096x2048_rgb_gather_{how}"] = sweep_bench("time_of_day", 4096, 2048, 8, how)
                sw[f"animation_128x8192x4096_spectral_gather_{how}"] = sweep_bench("animation", 8192, 4096, 2, how)
        for c_ in comm_box:
            c_.close()
        L.ssk_sweep_release()
        sw["note"] = ("BASELINE configs[3] (64 sun elevations x 4 turbidities, RGB) and configs[4] (128-frame turbidity/albedo "
                      "animation, spectral, wavelengths 440/520/600/680 nm); wall clock between barriers, max over ranks, includes "
                      "setting up every frame's emitter (host interpolation + batched device quadrature). gather_peer: every rank's "
                      "envmap kernels store straight into rank 0's IPC-mapped buffer over NVLink (no staging, no copy afterwards); "
                      "gather_nccl: grouped ncclSend/ncclRecv on a second stream, double-buffered against the next chunk's kernels")
        sweeps = sw

    if not args.no_extra and not args.sweeps_only:
        extra["note"] = ("configs[1] inputs for eval/pdf (2^24 directions), configs[2] for sample_direction (2^26 samples, "
                         "t=6 a=0.5 eta=50.2deg; sky only = sampling_test.py). fp32_frac_algorithmic = rate x the reference's flop count per unit (SURVEY 8d) / fp32 peak.")

    # ---- CPU baseline: the reference's own CPU code on the box's host cores (rank 0, N = 1 only) -------
    cpu = None
    if rank == 0 and world == 1 and not args.no_extra and not args.sweeps_only:
        try:
            import ref_oracle
            if ref_oracle.available():
                cores = os.cpu_count() or 1
                threads = min(cores, 64)
                wi_h, wav_h = wi.cpu().numpy(), wav.cpu().numpy()
                fl = cpu_flavours(wi_h, wav_h, threads)
                best = max(("packet", "scalar"), key=lambda k: fl[k]["value"])
                cpu = {"value": fl[best]["value"], "unit": "Gevals/s", "cores": threads, "kind": "reference",
                       "sample": f"{fl[best]['lanes']} of the {N_EVAL} configs[1] directions, best of >= 3 s of repeated passes ({fl[best]['seconds']:.2f} s each), the "
                                 f"reference's eval() on {'dr::Packet<float, %d> (%s)' % (fl['packet_width'], fl['isa']) if best == 'packet' else 'float'} "
                                 f"over {threads} host threads (the faster of its two CPU flavours; host has {cores} logical cores)",
                       "scalar_value": fl["scalar"]["value"], "scalar_single_thread_value": fl["scalar"]["single_thread_value"],
                       "packet_value": fl["packet"]["value"], "packet_single_thread_value": fl["packet"]["single_thread_value"],
                       "isa": fl["isa"], "packet_width": fl["packet_width"]}
                # BASELINE configs[0]: scalar_rgb eval over the 512 x 256 lat-long grid of test_sunsky.py:93-100 (t=3, a=0.3,
                # sun elevation 30 deg), one host thread; and the same grid on the GPU, compared with the reference's output
                grid = common.latlong_grid(512, 256)
                r0 = ref_oracle.RefEmitter(spectral=False, flavour=0, **kw)
                r0.eval(grid[:, :1024].copy())
                t_best = 1e30
                for _ in range(3):
                    t0 = time.perf_counter(); ref0 = r0.eval(grid); t_best = min(t_best, time.perf_counter() - t0)
                e0 = ssk.SunskyEmitter(spectral=False, device=local, arith="exact", **kw)
                g_t = torch.tensor(grid, device=dev)
                o_t = torch.empty((3, grid.shape[1]), device=dev)
                g_pp, o_pp = planes(g_t), planes(o_t)

                def step_grid():
                    assert L.ssk_eval_device(e0._ctx, g_pp, None, grid.shape[1], o_pp, stream) == 0
                ms_g, _ = timed(step_grid, 50, 5)
                ref1 = ref_oracle.RefEmitter(spectral=False, flavour=1, **kw).eval(grid)
                got = o_t.cpu().numpy()
                nz = ref0 != 0
                extra["configs0_scalar_rgb_latlong_512x256"] = {
                    "rays": int(grid.shape[1]),
                    "reference_scalar_rgb_1_thread_ms": t_best * 1e3, "reference_Mevals_s": grid.shape[1] / t_best / 1e6,
                    "gpu_exact_ms": ms_g / 50, "gpu_Mevals_s": grid.shape[1] / (ms_g / 50 * 1e-3) / 1e6,
                    "gpu_bit_identical_to_array_flavour": bool(np.array_equal(got.view(np.uint32), ref1.view(np.uint32))),
                    "gpu_vs_scalar_flavour_max_rel": float((np.abs(got[nz] - ref0[nz]) / np.abs(ref0[nz])).max()),
                    "note": "the GPU figure is launch-latency bound (131072 rays = 0.4 waves); parity is against both reference flavours"}
                e0.close()
        except Exception as ex:  # the baseline is reported, never required
            cpu = {"error": repr(ex)}

    sampler.stop()
    clocks = sampler.summary(headline_window)
    if clocks.get("window", "").startswith("whole"):
        clocks = sampler.summary(windows)
    if rank == 0:
        print(json.dumps({
            "metric": METRIC, "value": value, "unit": "Gevals/s", "n_gpus": world, "steps": args.steps,
            "warmup": warm, "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "f32", "data": "synthetic (PCG32 streams, SURVEY.md 8d cfg 2)",
            "config": CONFIG,
            "arith": args.arith + (" (MUFU transcendentals for well-conditioned terms, within 1e-5 of the reference; "
                                   "directions bit-identical)" if args.arith == "fast" else
                                   " (reference op order, bit-identical results)"),
            "roofline": roofline, "cpu_baseline": cpu, "e2e": e2e, "gpu_launches": launches, "clocks": clocks,
            "sweeps": sweeps, "extra": extra,
        }))
    if world > 1:
        dist.destroy_process_group()
    return 0


if __name__ == "__main__":
    # stdout carries exactly one JSON line: native libraries that write to fd 1 (NCCL prints its version banner
    # there) are diverted to stderr for the whole run, and the line goes out through the saved descriptor.
    sys.stdout.flush()
    _saved = os.dup(1)
    os.dup2(2, 1)
    sys.stdout = os.fdopen(_saved, "w")
    rc = main()
    sys.stdout.flush()
    sys.exit(rc)
