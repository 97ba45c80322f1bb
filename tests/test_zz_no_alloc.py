"""rt_render_paper_device_noalloc (include/rt_b200.h): the Paper frame through the entry point that allocates nothing
(G-buffer in a static __device__ array) equals rt_render_device's frame byte for byte, on two calls in a row, and
frames beyond RT_NOALLOC_MAX_PIXELS are rejected.  Collected last on purpose (pytest -x): named test_zz_*."""
import numpy as np
import pytest

from conftest import scene_path

pytestmark = pytest.mark.gpu


def _frames(gpu, name):
    import torch
    sc = gpu.load_scene_from_json(scene_path(name))
    tr = gpu.Tracer(sc, gpu.RenderMode.Paper)
    stream = torch.cuda.current_stream().cuda_stream
    want = torch.zeros((sc.height, sc.width, 4), dtype=torch.uint8, device="cuda")
    tr.render_device(want.data_ptr(), stream=stream)
    torch.cuda.synchronize()
    return sc, tr, stream, want


@pytest.mark.parametrize("names", [("penguin_small", "quirks")])
def test_noalloc_paper_frame_equals_render_device_twice(gpu, names):
    import torch
    for call, name in enumerate(names + names[:1]):        # three calls in a row, two scenes: the scratch is reused dirty
        sc, tr, stream, want = _frames(gpu, name)
        assert sc.width * sc.height <= gpu.NOALLOC_MAX_PIXELS
        got = torch.full((sc.height, sc.width, 4), 7, dtype=torch.uint8, device="cuda")
        free_before = torch.cuda.mem_get_info()[0]            # driver-level: sees the library's cudaMalloc, not only torch's
        tr.render_paper_device_noalloc(got.data_ptr(), stream=stream)
        torch.cuda.synchronize()
        assert torch.cuda.mem_get_info()[0] >= free_before, "the no-alloc entry point took device memory"
        w, g = want.cpu().numpy(), got.cpu().numpy()
        assert (g[..., 3] == 255).all()
        assert np.array_equal(g, w), f"call {call} ({name}): {(g != w).any(axis=2).sum()} pixels differ (tolerance: none, paper mode is pixel-exact)"
        assert w[..., :3].min() < 255 and w[..., :3].max() > 0, "degenerate frame"


def test_noalloc_sharded_ranks_cover_the_frame(gpu):
    import torch
    sc, tr, stream, want = _frames(gpu, "quirks")
    got = torch.zeros((sc.height, sc.width, 4), dtype=torch.uint8, device="cuda")
    for rank in range(3):
        tr.render_paper_device_noalloc(got.data_ptr(), stream=stream, rank=rank, world=3)
    torch.cuda.synchronize()
    assert torch.equal(got, want)


def test_noalloc_rejects_frames_beyond_the_limit(gpu):
    import torch
    sc = gpu.load_scene_from_json(scene_path("penguin_4k"))
    assert sc.width * sc.height > gpu.NOALLOC_MAX_PIXELS
    tr = gpu.Tracer(sc, gpu.RenderMode.Paper)
    buf = torch.zeros(16, dtype=torch.uint8, device="cuda")          # never written: the call must fail before any launch
    with pytest.raises(gpu.RtError) as e:
        tr.render_paper_device_noalloc(buf.data_ptr(), stream=torch.cuda.current_stream().cuda_stream)
    assert e.value.code == -4 and "RT_NOALLOC_MAX_PIXELS" in e.value.message
    assert not buf.any()
