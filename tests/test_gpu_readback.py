"""Row f3 of SURVEY 8(f): the library's asynchronous read-back pipeline (rt_render_to_host_async / rt_readback_async /
rt_wait_slot, csrc/api.cu) and the frame-sequence dump on top of it, through the C-ABI."""
import ctypes as C

import numpy as np
import pytest

import gpu_util as gu
import raytracing_cuda_b200 as rt
from raytracing_cuda_b200 import _lib, configs
from raytracing_cuda_b200.structs import Camera, Float3, Light, Object

pytestmark = pytest.mark.gpu


def test_async_slots_deliver_the_synchronous_frames():
    """Eight different frames through three slots, each read only after rt_wait_slot; slots are reused while earlier
    copies are still in flight (the staging frame's rewrite waits for its copy on the device, not on the host)."""
    r = gu.renderer()
    s = rt.Scene()
    w, h = 640, 360
    states = []
    for k in range(8):
        f = s.configure(view=5 + (k & 1), hours=3.0 * k, imgw=w, imgh=h, alias=bool(k % 3))
        states.append(configs.state(s, f))
    want = []
    for objs, lights, f in states:
        r.set_scene(objs, lights)
        want.append(r.render_to_host(f).copy())
    pinned = rt.PinnedFrames(3, w * h)
    got = [None] * len(states)
    for k, (objs, lights, f) in enumerate(states):
        if k >= 3:
            r.wait_slot(k % 3)
            got[k - 3] = pinned[k % 3].reshape(h, w).copy()
        r.set_scene(objs, lights)
        r.render_to_host_async(f, pinned[k % 3], k % 3)
    for k in range(len(states) - 3, len(states)):
        r.wait_slot(k % 3)
        got[k] = pinned[k % 3].reshape(h, w).copy()
    for k in range(len(states)):
        assert (got[k] == want[k]).all(), k
    # a sharded frame: pixels of other shards arrive as zeros, like rt_render_to_host
    sh = rt.make_shard(2, 1, 64, 32)
    objs, lights, f = states[1]
    r.set_scene(objs, lights)
    a = r.render_to_host(f, shard=sh, out=np.zeros(w * h, dtype=np.uint32)).copy()
    r.render_to_host_async(f, pinned[0], 0, shard=sh)
    r.wait_slot(0)
    assert (pinned[0].reshape(h, w) == a).all()
    with pytest.raises(_lib.RtError):
        r.render_to_host_async(f, pinned[0], _lib.RT_MAX_SLOTS)
    pinned.close()


def test_sequence_dump_writes_every_frame(tmp_path):
    r = gu.renderer()
    states = [(o, l, f) for _, o, l, f in configs.config5(frames=(0, 100, 200, 255), imgw=320, imgh=180)]
    seen = {}
    n = r.render_sequence(states, str(tmp_path / "path_"), fmt="png", depth=2,
                          on_frame=lambda i, img: seen.__setitem__(i, img.copy()))
    assert n == 4 and sorted(seen) == [0, 1, 2, 3]
    for i, (objs, lights, f) in enumerate(states):
        r.set_scene(objs, lights)
        want = r.render_to_host(f)
        assert (seen[i] == want).all()
        rgba = rt.read_png(str(tmp_path / ("path_%06d.png" % i)))
        assert (rgba[..., 0] == (want & 255)).all() and (rgba[..., 1] == ((want >> 8) & 255)).all() and \
            (rgba[..., 2] == ((want >> 16) & 255)).all()


def test_legacy_context_reads_back_what_launchKernel_rendered():
    """launchKernel leaves the frame in the caller's DEVICE buffer (reference kernel.cu:406-411); rt_readback_async on
    the context behind it (rt_legacy_context) queues the copy behind that frame without a host synchronisation."""
    import torch
    lib = _lib.load()
    s = rt.Scene()
    sky = gu.sky_small()
    w, h = 480, 270
    dev = [torch.zeros(w * h, dtype=torch.int32, device="cuda") for _ in range(2)]
    pinned = rt.PinnedFrames(2, w * h)
    want = []
    for k, hours in enumerate((14.0, 1.0)):
        f = s.configure(view=5, hours=hours, imgw=w, imgh=h, alias=True)
        cam = Camera()
        C.memmove(C.byref(cam), C.byref(f.cam), C.sizeof(Camera))
        lib.launchKernel(dev[k].data_ptr(), w, h, cam, C.cast(s.objects.ctypes.data, C.POINTER(Object)),
                         C.cast(s.lights.ctypes.data, C.POINTER(Light)), Float3(*f.ambient), (C.c_float * 4)(*f.skyVars),
                         sky[0].ctypes.data, sky[1].ctypes.data, sky[2].ctypes.data, sky[3].ctypes.data, sky.shape[2],
                         sky.shape[1], f.dayTime, True)
        rt.Renderer.legacy().readback_async(dev[k], w * h, pinned[k], k)
        r = gu.renderer()
        r.set_scene(s.objects, s.lights)
        want.append(r.render_to_host(f).copy())
    leg = rt.Renderer.legacy()
    for k in range(2):
        leg.wait_slot(k)
        assert (pinned[k].reshape(h, w) == want[k]).all(), k
    pinned.close()
