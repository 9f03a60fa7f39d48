"""rt_render_block_noalloc (include/rt_b200.h; no counterpart in the reference):
the block kernel with its scratch in a file-scope device array, for callers that must not allocate.

Collected LAST on purpose.  Passes alone in a fresh process and after the other
tests: it uses its own context and device buffers and only the C ABI.
"""
import ctypes as C

import numpy as np
import pytest

import openglraytracer_b200 as rt
from openglraytracer_b200 import _abi, api
from parity import assert_parity

pytestmark = pytest.mark.gpu


class _DeviceFrame:
    """A caller-owned device buffer through the C ABI (rt_device_alloc): allocated BEFORE the calls under test."""

    def __init__(self, shape):
        self.shape = tuple(shape)
        self.nbytes = int(np.prod(self.shape))
        self.ptr = C.c_void_p()
        _abi.check(_abi.lib().rt_device_alloc(C.byref(self.ptr), self.nbytes), "rt_device_alloc")

    def fetch(self):
        import torch
        view = torch.as_tensor(rt.multigpu._RawCudaBuffer(self.ptr.value, self.shape), device="cuda:0")
        return view.cpu().numpy().copy()

    def fill(self, value):
        import torch
        torch.as_tensor(rt.multigpu._RawCudaBuffer(self.ptr.value, self.shape), device="cuda:0").fill_(value)
        torch.cuda.synchronize()

    def free(self):
        _abi.lib().rt_device_free(self.ptr)


def _free_bytes():
    import torch
    torch.cuda.synchronize()
    return torch.cuda.mem_get_info(0)[0]


def test_no_alloc_entry_point_matches_the_existing_one():
    import openglraytracer_b200.multigpu  # noqa: F401  (the raw-buffer view used by _DeviceFrame)
    w, h = 384, 216
    with rt.Raytracer(0) as tr:
        new, old = _DeviceFrame((h, w, 4)), _DeviceFrame((h, w, 4))
        try:
            # two calls in a row on one scene (different cameras: the second call finds the first call's lists, counter
            # and camera in the shared array), then a different scene / depth / spp, then a 3-camera batch and a band
            cases = [(rt.scaled_scene(300, max_depth=0), rt.builtin_camera(0.0), 1),
                     (None, rt.builtin_camera(1.3), 1),
                     (rt.scaled_scene(1500, max_depth=1), rt.builtin_camera(0.4), 1),
                     (None, rt.builtin_camera(0.4), 4),
                     (rt.scaled_scene(128, max_depth=2), rt.builtin_camera(2.0), 1)]
            free_before = None
            for k, (sc, cam, spp) in enumerate(cases):
                if sc is not None:
                    tr.upload_scene(sc)                       # set-up: may allocate
                tr.render(None, cam, w, h, spp=spp, out_device_ptr=old.ptr.value)         # the existing entry point
                new.fill(0x5A)
                if k == 1:
                    free_before = _free_bytes()               # call 0 has loaded every kernel (lazy module loading)
                api.render_block_noalloc(tr, cam, w, h, new.ptr.value, spp=spp)
                if k == 1:
                    assert _free_bytes() == free_before, "the no-alloc call changed the device's free memory"
                a, b = new.fetch(), old.fetch()
                assert_parity(a, b, "no-alloc call %d" % k)  # the tolerance of the parity tests ...
                assert np.array_equal(a, b), "call %d"       # ... and in fact the same bytes: same kernel, same records
                assert int(a[..., 3].max()) == 0
            # batch of 3 cameras + a row band written in place
            cams = rt.turntable_cameras(3)
            nb, ob_ = _DeviceFrame((3, h, w, 4)), _DeviceFrame((3, h, w, 4))
            tr.render(None, cams, w, h, out_device_ptr=ob_.ptr.value)
            api.render_block_noalloc(tr, cams, w, h, nb.ptr.value)
            assert np.array_equal(nb.fetch(), ob_.fetch())
            new.fill(7)
            api.render_block_noalloc(tr, rt.builtin_camera(2.0), w, h, new.ptr.value, y0=40, y1=100)
            tr.render(None, rt.builtin_camera(2.0), w, h, out_device_ptr=old.ptr.value)
            got, want = new.fetch(), old.fetch()
            assert np.array_equal(got[40:100], want[40:100]) and (got[:40] == 7).all() and (got[100:] == 7).all()
            nb.free(); ob_.free()

            # limits: rejected with a clear message, nothing rendered
            tr.upload_scene(rt.scaled_scene(_abi.RT_NOALLOC_MAX_OBJECTS + 1))
            with pytest.raises(rt.RtError) as ei:
                api.render_block_noalloc(tr, rt.builtin_camera(0.0), w, h, new.ptr.value)
            assert ei.value.code == _abi.RT_ERR_UNSUPPORTED and "RT_NOALLOC_MAX_OBJECTS" in str(ei.value)
            tr.upload_scene(rt.scaled_scene(_abi.RT_NOALLOC_MAX_OBJECTS))
            api.render_block_noalloc(tr, rt.builtin_camera(0.0), w, h, new.ptr.value)      # the stated maximum works
            tr.render(None, rt.builtin_camera(0.0), w, h, out_device_ptr=old.ptr.value)
            assert np.array_equal(new.fetch(), old.fetch())
            with pytest.raises(rt.RtError) as ei:
                api.render_block_noalloc(tr, rt.turntable_cameras(_abi.RT_NOALLOC_MAX_FRAMES + 1), 64, 36, new.ptr.value)
            assert ei.value.code == _abi.RT_ERR_UNSUPPORTED and "RT_NOALLOC_MAX_FRAMES" in str(ei.value)
            tr.upload_scene(rt.builtin_scene(0.0, 0))
            with pytest.raises(rt.RtError) as ei:
                api.render_block_noalloc(tr, rt.builtin_camera(0.0), w, h, new.ptr.value)
            assert ei.value.code == _abi.RT_ERR_UNSUPPORTED
        finally:
            new.free(); old.free()
