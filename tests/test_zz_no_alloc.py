"""mcpt_render_accumulate_device_noalloc (include/mcpt.h; csrc/noalloc.cu): the render + slice reduce with its scratch
in a fixed file-scope __device__ array instead of cudaMalloc'ed buffers.  It has no counterpart in the reference,
so the check is against the library's own allocating entry point
mcpt_render_accumulate_device through the C ABI: same bits where the slice count is the same, fp32 reassociation where the
fixed scratch holds fewer slices.  Two calls in a row (the array keeps the previous call's contents), device memory
unchanged across a call, limits rejected with an error.  Collected last on purpose (pytest -x)."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu

W = H = 128      # 16 384 pixels x 37 slices = 606 208 entries: fits the 2**20-entry scratch with the allocating path's slice count
SPP, DEPTH = 64, 8


def _accumulate(fn, T, cfgs, w, h):
    import torch
    acc = torch.zeros(h, w, 4, device="cuda", dtype=torch.float32)
    stream = torch.cuda.current_stream().cuda_stream
    torch.cuda.synchronize()      # stream 0 = the library's own non-blocking stream: the zero fill must have landed
    for cfg in cfgs:
        fn(cfg, acc, stream)
    torch.cuda.synchronize()
    return acc.cpu().numpy()


def test_noalloc_render_matches_allocating_entry_point(gpu, scenes):
    import torch
    T, s = gpu, scenes["scene01"]
    T.create_geometry(s.verts, s.normals, s.tris, s.mats, T.BUILD_SAH)
    alloc = lambda cfg, acc, st: T.render_accumulate_device(cfg, acc.data_ptr(), 0, st)
    noalloc = lambda cfg, acc, st: T.render_accumulate_device_noalloc(cfg, acc.data_ptr(), st)

    # two calls in a row, different sample ranges and seeds: the second call finds the first call's slices and counters
    cfgs = [T.render_cfg(W, H, SPP, DEPTH, seed=7), T.render_cfg(W, H, SPP // 2, DEPTH, seed=7, sample_offset=SPP)]
    want1 = _accumulate(alloc, T, cfgs[:1], W, H)
    rays1 = T.counters().rays
    want2 = _accumulate(alloc, T, cfgs, W, H)
    got1 = _accumulate(noalloc, T, cfgs[:1], W, H)
    c = T.counters()
    assert c.rays == rays1 and c.paths == W * H * SPP and c.launches == 2
    got2 = _accumulate(noalloc, T, cfgs, W, H)
    assert (got1[..., 3] == SPP).all() and (got2[..., 3] == SPP + SPP // 2).all()
    assert np.array_equal(got1, want1), "first no-alloc call differs from mcpt_render_accumulate_device"
    assert np.array_equal(got2, want2), "second no-alloc call differs (stale scratch?)"
    assert got1[..., :3].max() > 0

    # a smaller image right after a larger one (stale entries beyond its range), with next-event estimation
    cfg_small = T.render_cfg(48, 40, 16, 5, seed=3, flags=T.FLAG_NEE)
    assert np.array_equal(_accumulate(noalloc, T, [cfg_small], 48, 40), _accumulate(alloc, T, [cfg_small], 48, 40))

    # more pixels than the scratch holds slices for at the allocating path's slice count: fewer slices, same samples,
    # other fp32 summation order (tolerance of test_accumulate_device_equals_one_call: 1e-6 on the mean, relative here)
    cfg_big = T.render_cfg(768, 768, 128, DEPTH, seed=5)
    a = _accumulate(alloc, T, [cfg_big], 768, 768)
    b = _accumulate(noalloc, T, [cfg_big], 768, 768)
    assert (b[..., 3] == 128).all()
    assert np.allclose(b[..., :3] / 128, a[..., :3] / 128, rtol=1e-5, atol=1e-6)

    # no device allocation or release across a call (the module and its array are loaded by now)
    acc = torch.zeros(H, W, 4, device="cuda", dtype=torch.float32)
    torch.cuda.synchronize()
    free1, _ = torch.cuda.mem_get_info()
    T.render_accumulate_device_noalloc(cfgs[0], acc.data_ptr(), torch.cuda.current_stream().cuda_stream)
    free2, _ = torch.cuda.mem_get_info()
    assert free2 == free1, (free1, free2)
    assert np.array_equal(acc.cpu().numpy(), want1)


def test_noalloc_render_rejects_what_it_cannot_hold(gpu, scenes):
    import torch
    T, s = gpu, scenes["scene01"]
    T.create_geometry(s.verts, s.normals, s.tris, s.mats, T.BUILD_SAH)
    acc = torch.zeros(1025, 1024, 4, device="cuda", dtype=torch.float32)
    torch.cuda.synchronize()
    with pytest.raises(RuntimeError, match="MCPT_NOALLOC_MAX_SLICE_PIXELS"):
        T.render_accumulate_device_noalloc(T.render_cfg(1024, 1025, 1, 5), acc.data_ptr())
    assert (acc == 0).all()
    with pytest.raises(RuntimeError, match="PHILOX"):
        T.render_accumulate_device_noalloc(T.render_cfg(64, 64, 1, 5, rng_mode=T.RNG_REFERENCE), acc.data_ptr())
    with pytest.raises(RuntimeError, match="MOMENTS"):
        cfg = T.render_cfg(64, 64, 1, 5)
        cfg.flags |= T.FLAG_MOMENTS
        T.render_accumulate_device_noalloc(cfg, acc.data_ptr())
    with pytest.raises(RuntimeError, match="null dev_sum"):
        T.render_accumulate_device_noalloc(T.render_cfg(64, 64, 1, 5), 0)
    T.destroy_geometry()
    with pytest.raises(RuntimeError, match="no geometry resident"):
        T.render_accumulate_device_noalloc(T.render_cfg(64, 64, 1, 5), acc.data_ptr())
    # the largest size it takes: exactly 2**20 pixels, one slice
    T.create_geometry(s.verts, s.normals, s.tris, s.mats, T.BUILD_SAH)
    acc = torch.zeros(1024, 1024, 4, device="cuda", dtype=torch.float32)
    ref = torch.zeros(1024, 1024, 4, device="cuda", dtype=torch.float32)
    torch.cuda.synchronize()
    T.render_accumulate_device_noalloc(T.render_cfg(1024, 1024, 2, 5, seed=1), acc.data_ptr())
    T.render_accumulate_device(T.render_cfg(1024, 1024, 2, 5, seed=1), ref.data_ptr(), 0, 0)
    assert torch.equal(acc, ref)
