"""GPU box: the reference's micro-benchmark list (gpu/bench.cu:435-451) re-run like for like.
Left column: the reference's own functions compiled unmodified for sm_100
(oracle/_ref/libref_gpu.so: parallel_radix_sort, shared_prefix_sum, to_cuda = upload +
partitioning build).  Right column: librtb200's device primitives and rtb_scene_create on
the same inputs.  Results are checked against numpy before a time is printed.
usage: python tests/reports/microbench.py [--sizes 100000,1000000,10000000] [--ref-max 10000000]"""
import argparse
import ctypes as C
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(ROOT, "raytracing-gpu_b200"))
sys.path.insert(0, os.path.join(ROOT, "tests"))  # util.py
import pyrtb  # noqa: E402
import torch  # noqa: E402  (device buffers + events only)

ap = argparse.ArgumentParser()
ap.add_argument("--sizes", default="100000,1000000,10000000")
ap.add_argument("--ref-max", type=int, default=10000000)
ap.add_argument("--iters", type=int, default=5)
a = ap.parse_args()

R = C.CDLL(os.path.join(ROOT, "oracle", "_ref", "libref_gpu.so"))
R.refgpu_bench_sort_u32.argtypes = [C.c_void_p, C.c_size_t, C.c_int, C.POINTER(C.c_double), C.c_void_p]
R.refgpu_bench_scan_u64.argtypes = [C.c_void_p, C.c_size_t, C.c_int, C.POINTER(C.c_double), C.c_void_p]
R.refgpu_bench_to_cuda.argtypes = [C.c_void_p, C.c_int, C.POINTER(C.c_double)]
L = pyrtb.lib()
L.rtb_dev_radix_sort_pairs_u64.argtypes = [C.c_void_p, C.c_void_p, C.c_size_t, C.c_int, C.c_void_p]
L.rtb_dev_inclusive_scan_u32.argtypes = [C.c_void_p, C.c_size_t, C.c_void_p]
dev = torch.device("cuda:0")
stream = torch.cuda.current_stream(dev)
sp = C.c_void_p(stream.cuda_stream)


def timed(fn, prep, iters):
    best = 1e30
    for _ in range(iters):
        prep()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        torch.cuda.synchronize()
        e0.record(stream)
        fn()
        e1.record(stream)
        torch.cuda.synchronize()
        best = min(best, e0.elapsed_time(e1))
    return best


print("%-34s %10s %14s %14s %9s" % ("primitive", "n", "reference ms", "librtb200 ms", "ratio"))
rng = np.random.default_rng(0xB200)
for n in [int(x) for x in a.sizes.split(",")]:
    # --- radix sort of u32 keys with a payload (bench.cu:224-257: size_t + uint16_t payloads) ---
    keys = rng.integers(0, 2**31, n, dtype=np.uint32)       # rand() range, as bench.cu:231
    want = np.sort(keys, kind="stable")
    ref_ms = None
    if n <= a.ref_max:
        out = np.zeros(n, np.uint32)
        ms = C.c_double()
        rc = R.refgpu_bench_sort_u32(keys.ctypes.data, n, max(1, a.iters // 2), C.byref(ms), out.ctypes.data)
        assert rc == 0 and np.array_equal(out, want), "reference sort failed (rc %d)" % rc
        ref_ms = ms.value
    h64 = torch.from_numpy(keys.astype(np.uint64).view(np.int64))
    dk = torch.empty(n, dtype=torch.int64, device=dev)
    dv = torch.empty(n, dtype=torch.int32, device=dev)
    iota = torch.arange(n, dtype=torch.int32, device=dev)

    def prep():
        dk.copy_(h64)
        dv.copy_(iota)

    def run():
        rc = L.rtb_dev_radix_sort_pairs_u64(dk.data_ptr(), dv.data_ptr(), n, 32, sp)
        pyrtb._check(rc)
    ours = timed(run, prep, a.iters)
    got = dk.cpu().numpy().view(np.uint64)
    perm = dv.cpu().numpy()
    assert np.array_equal(got, want.astype(np.uint64)) and np.array_equal(keys[perm], want), "librtb200 sort wrong"
    assert np.array_equal(perm, np.argsort(keys, kind="stable").astype(np.int32)), "librtb200 sort not stable"
    print("%-34s %10d %14s %14.3f %9s" % ("radix sort u32 keys + payload", n, "%.3f" % ref_ms if ref_ms else "-", ours,
                                          "%.0fx" % (ref_ms / ours) if ref_ms else "-"), flush=True)

    # --- inclusive prefix sum (bench.cu BM_prefix_sum_parallel: size_t values) ---
    vals = rng.integers(0, 9, n, dtype=np.uint64)
    want = np.cumsum(vals, dtype=np.uint64)
    ref_ms = None
    if n <= a.ref_max:
        out = np.zeros(n, np.uint64)
        ms = C.c_double()
        rc = R.refgpu_bench_scan_u64(vals.ctypes.data, n, a.iters, C.byref(ms), out.ctypes.data)
        assert rc == 0 and np.array_equal(out, want), "reference scan failed (rc %d)" % rc
        ref_ms = ms.value
    h32 = torch.from_numpy(vals.astype(np.uint32).view(np.int32))
    dd = torch.empty(n, dtype=torch.int32, device=dev)

    def prep2():
        dd.copy_(h32)

    def run2():
        rc = L.rtb_dev_inclusive_scan_u32(dd.data_ptr(), n, sp)
        pyrtb._check(rc)
    ours = timed(run2, prep2, a.iters)
    assert np.array_equal(dd.cpu().numpy().view(np.uint32).astype(np.uint64), want), "librtb200 scan wrong"
    print("%-34s %10d %14s %14.3f %9s" % ("inclusive scan (ref u64, ours u32)", n, "%.3f" % ref_ms if ref_ms else "-",
                                          ours, "%.0fx" % (ref_ms / ours) if ref_ms else "-"), flush=True)

# --- upload + partitioning build on the scenes of FULL_BENCHMARK (bench.cu:18-22) ---
from util import load_golden  # noqa: E402
print()
print("%-34s %10s %14s %14s %9s" % ("upload + partitioning build", "triangles", "reference ms", "librtb200 ms", "ratio"))
for name in ("cube", "island_smooth", "dark-night", "spheres"):
    hs, _ = load_golden(name)
    s = hs.scaled(3)
    ms = C.c_double()
    rc = R.refgpu_bench_to_cuda(C.byref(s), a.iters, C.byref(ms))
    assert rc == 0
    best = 1e30
    for _ in range(a.iters + 1):
        t0 = time.perf_counter()
        ds = pyrtb.DeviceScene(hs, 3)
        best = min(best, (time.perf_counter() - t0) * 1e3)
        st = ds.stats()
        ds.close()
    print("%-34s %10d %14.3f %14.3f %8.1fx   (device part of ours: upload %.3f + build %.3f ms, %d nodes)" % (
        name, hs.triangle_count, ms.value, best, ms.value / best, st["upload_ms"], st["build_ms"], st["nodes"]), flush=True)
