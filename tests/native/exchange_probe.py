"""Body of tests/test_gpu_raster.py::test_sharded_projection_pass_with_item_exchange (GPU box): N scene
handles on one device act as the N ranks of a sharded frame.  Three frames: both halves of the
double-buffered inboxes and an epoch that comes back to a parity.  The composed frame must be the
single-rank frame byte for byte, nobody may give up or time out.
usage: CUDA_DEVICE_MAX_CONNECTIONS=32 python tests/native/exchange_probe.py N"""
import ctypes as C
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(ROOT, "raytracing-gpu_b200"))
import pyrtb  # noqa: E402

nranks = int(sys.argv[1])
L = pyrtb.lib()
hs = pyrtb.HostScene.terrain(250, 640, 400)          # 125 000 triangles
ranks = [pyrtb.DeviceScene(hs, 3) for _ in range(nranks)]
ranks[0].primary = 2
want = ranks[0].render()
ranks[0].primary = 0
d_frame = C.c_void_p()
nbytes = 640 * 400 * 4
pyrtb._check(L.rtb_dev_alloc(0, nbytes, C.byref(d_frame)))
# one replicated frame per rank first: every handle allocates its buffers now (a host-side allocation
# synchronises the device, which on ONE device would wait for another rank's spinning barrier kernel)
for r, h in enumerate(ranks):
    h.render_device(d_frame, 640 * 4, rank=r, nranks=nranks)
    h.sync()
blocks = [h.exchange_create(nranks)[0] for h in ranks]
for r, h in enumerate(ranks):
    h.exchange_attach(r, nranks, blocks)
for frame in range(3):
    pyrtb._check(L.rtb_dev_memset(d_frame, 0, nbytes))
    for r, h in enumerate(ranks):
        h.render_device(d_frame, 640 * 4, rank=r, nranks=nranks)       # asynchronous: nobody waits on the host
    for h in ranks:
        h.sync()
    got = np.zeros((400, 640, 4), np.uint8)
    pyrtb._check(L.rtb_dev_download(got.ctypes.data, d_frame, nbytes))
    for h in ranks:
        st = h.raster_state()
        assert not st["gave_up"], (frame, st)
        assert h.stats()["kernel_launches"] == 5, "clear/setup + big + exchange barrier + items + frame"
    assert sum(h.raster_state()["items"] for h in ranks) > 50000
    assert np.array_equal(got, want), (nranks, frame, int((got != want).any(axis=-1).sum()))
L.rtb_dev_free(d_frame)
for h in ranks:
    h.close()
print("exchange ok: %d ranks, 3 frames" % nranks)
