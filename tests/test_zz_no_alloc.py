"""zb_render_device_noalloc (csrc/noalloc.cu; no counterpart in the reference):
the beauty pass with its scratch in a static __device__ array instead of a cudaMalloc'ed buffer.  It runs the same kernels
on the same Philox keys as zb_render_device, so the sums must be EQUAL bit for bit — on two calls in a row (the second on
another frame size and seed, so stale scratch contents would show), under a sample slice with accumulation, and a frame
beyond the static maximum must be refused.  Collected last (file name) so that a failure here hides no other test; passes
alone in a fresh process and after the other tests (own context, own scene)."""
import numpy as np
import pytest

from raytracer_project_b200 import api, scenes

pytestmark = pytest.mark.gpu


def test_noalloc_entry_point_equals_render_device():
    import torch
    txt, _ = scenes.weekend(seed=1)
    flat = api.flatten_zscene(txt)
    with api.Context(0) as ctx:
        sc = ctx.upload(flat)
        api.free_flat(flat)
        # all buffers exist before the calls under test: nothing below allocates on the device
        bufs = {}
        for w, h in ((203, 90), (160, 96)):
            bufs[(w, h)] = (torch.full((w * h, 4), -1.0, device="cuda"), torch.full((w * h, 4), -2.0, device="cuda"))
        torch.cuda.synchronize()
        for call, ((w, h), spp, seed) in enumerate((((203, 90), 24, 5), ((160, 96), 16, 9))):
            cfg = scenes.weekend(seed=1, width=w, height=h, spp=spp)[1]
            a, b = bufs[(w, h)]
            sc.render_device(cfg, a.data_ptr(), seed=seed)
            sc.render_device_noalloc(cfg, b.data_ptr(), seed=seed)
            torch.cuda.synchronize()
            assert torch.equal(a, b), (call, float((a - b).abs().max()))
            assert float(a[:, 3].min()) == spp and float(a[:, :3].mean()) > 0.1
            # second half of the samples added to the first half: slices + accumulate through the static scratch
            sc.render_device_noalloc(cfg, b.data_ptr(), seed=seed, sample_range=(0, spp // 2))
            sc.render_device_noalloc(cfg, b.data_ptr(), seed=seed, sample_range=(spp // 2, spp), accumulate=True)
            torch.cuda.synchronize()
            assert np.allclose(a.cpu().numpy(), b.cpu().numpy(), rtol=2e-6, atol=1e-7), call
        # beyond the static maximum: refused with a clear error, nothing launched
        big = scenes.weekend(seed=1, width=2048, height=1024, spp=1)[1]
        assert 2048 * 1024 > ctx.lib.zb_noalloc_max_pixels() - 1
        big.image_height = 1025
        with pytest.raises(api._lib.ZbError) as e:
            sc.render_device_noalloc(big, bufs[(203, 90)][0].data_ptr())
        assert "ZB_NOALLOC_MAX_PIXELS" in str(e.value)
        sc.close()
