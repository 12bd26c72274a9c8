"""Measures the roofline denominators MEASURED_PEAKS.json does not carry on the GPU it runs
on and prints / writes them as JSON: L2 read bandwidth and L2 reduction (red.global.add.v4.f32) bandwidth on
buffers that fit L2, aggregate L1 read bandwidth, and the dense tf32 tcgen05 peak.  bench.py imports `measure()` for
a short live measurement; `python scripts/measure_peaks.py profiles/peaks_l2_tf32.json` records the tracked copy.
Kernels: fewtensorf_b200/csrc/k_peaks.cu (libftrf_peaks.so, a measurement aid outside the product library)."""
import ctypes as C
import json
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = os.path.join(ROOT, "fewtensorf_b200", "libftrf_peaks.so")


def _lib():
    lib = C.CDLL(LIB)
    lib.ftrf_peak_l2_read.argtypes = [C.c_void_p, C.c_size_t, C.c_int, C.c_int, C.c_void_p, C.c_void_p]
    lib.ftrf_peak_l2_red.argtypes = [C.c_void_p, C.c_size_t, C.c_int, C.c_int, C.c_void_p]
    lib.ftrf_peak_l1_read.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_void_p, C.c_void_p]
    lib.ftrf_peak_tf32.argtypes = [C.c_int, C.c_int, C.c_int, C.c_void_p]
    return lib


def _time(fn, reps=5):
    fn()
    torch.cuda.synchronize()
    best = float("inf")
    for _ in range(reps):
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        fn()
        b.record()
        b.synchronize()
        best = min(best, a.elapsed_time(b))
    return best * 1e-3


def measure(dev=None, quick=False):
    """dict of measured peaks (GB/s, TFLOP/s) on the current device; `quick` shortens every loop (~60 ms in all)."""
    lib = _lib()
    dev = dev or torch.device("cuda", torch.cuda.current_device())
    st = lambda: torch.cuda.current_stream().cuda_stream
    sink = torch.zeros(4, device=dev)
    sms = torch.cuda.get_device_properties(dev).multi_processor_count
    out = {"gpu": torch.cuda.get_device_name(dev), "sms": sms, "l2_read": {}, "l2_red": {}}
    for mb in ((32,) if quick else (8, 16, 24, 32, 40, 64)):
        buf = torch.zeros(mb << 18, device=dev)                    # mb MiB of fp32
        n = buf.numel() * 4
        iters = max(4, ((1024 if quick else 4096) << 20) // n)      # same bytes per launch whatever the set: the launch ramp is amortised alike
        for ctas in (sms * 2, sms * 4):
            t = _time(lambda: lib.ftrf_peak_l2_read(buf.data_ptr(), n, iters, ctas, sink.data_ptr(), st()))
            out["l2_read"][f"{mb}MB_x{ctas}"] = round(n * iters / t / 1e9, 1)
            t = _time(lambda: lib.ftrf_peak_l2_red(buf.data_ptr(), n, iters, ctas, st()))
            out["l2_red"][f"{mb}MB_x{ctas}"] = round(n * iters / t / 1e9, 1)
        del buf
    buf = torch.zeros(sms * 4 * 8192, device=dev)
    it1 = 2000 if quick else 20000
    t = _time(lambda: lib.ftrf_peak_l1_read(buf.data_ptr(), it1, sms * 4, sink.data_ptr(), st()))
    out["l1_read_gbs"] = round(sms * 4 * it1 * 32768 / t / 1e9, 1)
    out["tf32"] = {}
    itm = 4000 if quick else 40000
    for mode, name in ((0, "ss_n256_noswizzle"), (1, "ss_n256_sw128")):
        t = _time(lambda: lib.ftrf_peak_tf32(mode, itm, sms, st()), reps=3)
        out["tf32"][name] = round(sms * itm * 2 * 128 * 256 * 8 / t / 1e12, 1)
    # a working set must fit L2 with room to spare: the peak is the best figure among the <= 40 MB sets
    fits = lambda d: max(v for k, v in d.items() if int(k.split("MB")[0]) <= 40)
    out["l2_read_gbs"] = fits(out["l2_read"])
    out["l2_red_gbs"] = fits(out["l2_red"])
    out["tf32_tflops"] = max(out["tf32"].values())
    out["how"] = ("l2_read: ld.global.cg.v4 sweeps of a resident buffer, 8 loads in flight per thread; l2_red: red.global.add.v4.f32 "
                  "sweeps of the same buffer; l1_read: ld.global.ca.v4 over a 32 KB slice per CTA; tf32: back-to-back "
                  "tcgen05.mma.kind::tf32 M128 N256 K8 from shared memory, one CTA per SM; best of 5 (3), CUDA events")
    return out


if __name__ == "__main__":
    res = measure()
    print(json.dumps(res, indent=1))
    if len(sys.argv) > 1:
        with open(sys.argv[1], "w") as fh:
            json.dump(res, fh, indent=1)
