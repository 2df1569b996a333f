"""vkr_pack_rgb8_static (include/vkr.h): the allocation-free variant of the RGB8 read-back — pack_rgb8_kernel with its packed
image staged in a file-scope static device array instead of a cudaMalloc'd buffer.  It has no counterpart in the reference,
so there is no oracle leg here: the check is against the existing entry
point, vkr_readback(VKR_TARGET_RGB8), byte for byte (integer work: the bar is bit-exact), on two calls in a row, because the
array is shared by every call and keeps the last call's bytes.

Collected last on purpose: it passes alone in a fresh process and after the other tests.
"""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _frames(vkr_lib, width, height, scene_names):
    from vk_renderer_b200 import capi
    import scenes as sc
    r = capi.Renderer(vkr_lib, "vkr", width, height, shadow_extent=256)
    for name in scene_names:
        r.draw(*sc.SCENES[name](r))
        want = r.readback(capi.TARGET_RGB8)                      # the existing entry point (synchronous: the frame is complete)
        ptr, nbytes = r.device_ptr(capi.TARGET_RGBA8)
        assert nbytes == width * height * 4
        yield r, ptr, want.reshape(-1)
    r.close()


def test_static_pack_matches_the_rgb8_readback_on_two_calls_in_a_row(vkr_lib):
    """Two different frames, one after the other, through the shared static array: each must equal vkr_readback(RGB8) of its
    own frame (nothing of the first call may show through in the second), and a smaller image after a larger one as well."""
    from vk_renderer_b200 import capi
    seen = []
    for w, h, scenes_ in ((320, 200, ("basic_pcf", "transparent")), (161, 67, ("basic_pcf",))):      # 161*67: pixel count not a multiple of 4
        for r, ptr, want in _frames(vkr_lib, w, h, scenes_):
            out = np.full(w * h * 3 + 5, 0xA5, np.uint8)                                                # canary bytes behind the image
            r.pack_rgb8_static(ptr, w, h, out)
            assert np.array_equal(out[: w * h * 3], want), (w, h)
            assert (out[w * h * 3:] == 0xA5).all()
            seen.append(want.copy())
    assert not np.array_equal(seen[0], seen[1])           # the two calls in a row really carried different images


def test_static_pack_repeats_and_runs_on_a_caller_stream(vkr_lib):
    """The same image twice gives the same bytes; wait=0 on a stream of the caller's, synchronised by the caller."""
    import torch
    from vk_renderer_b200 import capi
    w, h = 256, 128
    for r, ptr, want in _frames(vkr_lib, w, h, ("basic_pcf",)):
        a = r.pack_rgb8_static(ptr, w, h, np.zeros(w * h * 3, np.uint8))
        b = r.pack_rgb8_static(ptr, w, h, np.zeros(w * h * 3, np.uint8))
        assert np.array_equal(a, want) and np.array_equal(b, want)
        st = torch.cuda.Stream(device=0)
        pinned = torch.zeros(w * h * 3, dtype=torch.uint8).pin_memory()
        r.pack_rgb8_static(ptr, w, h, pinned.numpy(), stream=st.cuda_stream, wait=False)
        st.synchronize()
        assert np.array_equal(pinned.numpy(), want)


def test_static_pack_refuses_what_does_not_fit(vkr_lib):
    from vk_renderer_b200 import capi
    w, h = 64, 64
    for r, ptr, want in _frames(vkr_lib, w, h, ("basic_pcf",)):
        big_w, big_h = 2048, 1025                         # 3 * 2048 * 1025 bytes: one row over VKR_PACK_STATIC_MAX_BYTES
        assert big_w * big_h * 3 > capi.PACK_STATIC_MAX_BYTES >= 1920 * 1080 * 3
        with pytest.raises(capi.VkrError) as e:
            r.pack_rgb8_static(ptr, big_w, big_h, np.zeros(big_w * big_h * 3, np.uint8))
        assert e.value.code == capi.ERROR_CAPACITY and "VKR_PACK_STATIC_MAX_BYTES" in str(e.value)
        with pytest.raises(capi.VkrError):                # destination too small
            r.pack_rgb8_static(ptr, w, h, np.zeros(w * h * 3 - 1, np.uint8))
        with pytest.raises(capi.VkrError):                # null image
            r.pack_rgb8_static(0, w, h, np.zeros(w * h * 3, np.uint8))
        # ... and the entry point still works after the refusals
        assert np.array_equal(r.pack_rgb8_static(ptr, w, h, np.zeros(w * h * 3, np.uint8)), want)
