"""GPU tier, collected last on purpose: vkrt_trace_rays_static, the ray-level entry
that allocates nothing (include/vkrt.h; no counterpart in the reference).
Its scratch is one file-scope device array of the library, shared by every call on a GPU, so each case calls it at
least twice in a row and compares every call, bit for bit, with vkrt_trace_rays on the same rays."""
import numpy as np
import pytest

import helpers as H
from vkrt_b200 import binding as B
from vkrt_b200 import scenes
from vkrt_b200.raytracer import flatten_model

pytestmark = pytest.mark.gpu

MISS = 0xFFFFFFFF


def same_records(a, b):
    return a.dtype == b.dtype and a.shape == b.shape and a.tobytes() == b.tobytes()


def test_static_entry_matches_trace_rays_single_instance_two_calls_in_a_row():
    flat = flatten_model(scenes.random_soup(6000, seed=3))
    inst = scenes.identity_instance(1.0)
    ctx = B.Context(0)
    blas = B.Blas(ctx, flat["vertices"], flat["indices"], flat["ranges"])
    tlas = B.Tlas(ctx, [blas], inst)
    first = H.random_rays(B.STATIC_MAX_RAYS, 11, (-12, -8, -10), (12, 8, 10))  # exactly the maximum
    first["tmax"][::7] = 3.0
    first["direction"][5] = (0, 0, 0)  # rejected rays report a miss
    first["direction"][6] = (np.nan, 1, 0)
    second = H.random_rays(20001, 12, (-12, -8, -10), (12, 8, 10))  # shorter and ragged: the array still holds call one
    want = [tlas.trace_rays(first), tlas.trace_rays(second)]
    got = [tlas.trace_rays_static(first), tlas.trace_rays_static(second)]
    assert same_records(got[0], want[0]) and same_records(got[1], want[1])
    assert (want[0]["instance"] != MISS).sum() > 1000 and (want[0]["instance"] == MISS).sum() > 1000
    # the same batch again, and any-hit right after closest-hit
    assert same_records(tlas.trace_rays_static(second), want[1])
    for rays in (first, second):
        a, b = tlas.trace_rays(rays, any_hit=True), tlas.trace_rays_static(rays, any_hit=True)
        assert same_records(a, b)
    # the two entry points interleaved: neither disturbs the other
    assert same_records(tlas.trace_rays(first), want[0]) and same_records(tlas.trace_rays_static(first), want[0])
    tlas.close(), blas.close(), ctx.close()


def test_static_entry_matches_trace_rays_top_level_and_second_context():
    fa, fb = flatten_model(scenes.random_soup(2000, seed=1)), flatten_model(scenes.random_soup(700, seed=2))
    ctx = B.Context(0)
    a = B.Blas(ctx, fa["vertices"], fa["indices"], fa["ranges"])
    b = B.Blas(ctx, fb["vertices"], fb["indices"], fb["ranges"], flags=B.BUILD_PREFER_FAST_BUILD)
    inst = scenes.random_instance_transforms(40, seed=7, extent=(60.0, 20.0, 60.0), scale=(0.5, 1.2))
    tlas = B.Tlas(ctx, [a if i % 2 == 0 else b for i in range(40)], inst)
    for seed, n in ((8, 30000), (9, 777)):
        rays = H.random_rays(n, seed, (-40, -15, -40), (40, 15, 40))
        rays["tmax"][::5] = 25.0
        want = tlas.trace_rays(rays)
        assert same_records(tlas.trace_rays_static(rays), want)
        assert len(np.unique(want["instance"][want["instance"] != MISS])) > 10
        assert same_records(tlas.trace_rays_static(rays, any_hit=True), tlas.trace_rays(rays, any_hit=True))
    # a second context on the same GPU shares the array (calls are serial here, as the header requires)
    ctx2 = B.Context(0)
    flat = flatten_model(scenes.analytic_scene())
    blas2 = B.Blas(ctx2, flat["vertices"], flat["indices"], flat["ranges"])
    tlas2 = B.Tlas(ctx2, [blas2], scenes.identity_instance(1.0))
    rays2 = H.random_rays(5000, 4, (-6, 0.5, -6), (6, 6, 6))
    assert same_records(tlas2.trace_rays_static(rays2), tlas2.trace_rays(rays2))
    rays = H.random_rays(4096, 10, (-40, -15, -40), (40, 15, 40))
    assert same_records(tlas.trace_rays_static(rays), tlas.trace_rays(rays))
    tlas2.close(), blas2.close(), ctx2.close()
    tlas.close(), a.close(), b.close(), ctx.close()


def test_static_entry_limits_and_pointer_kinds():
    flat = flatten_model(scenes.random_soup(1500, seed=5))
    ctx = B.Context(0)
    blas = B.Blas(ctx, flat["vertices"], flat["indices"], flat["ranges"])
    tlas = B.Tlas(ctx, [blas], scenes.identity_instance(1.0))
    rays = H.random_rays(B.STATIC_MAX_RAYS + 1, 6, (-12, -8, -10), (12, 8, 10))
    with pytest.raises(B.VkrtError) as e:
        tlas.trace_rays_static(rays)
    assert e.value.code == 1 and "at most 65536 rays" in str(e.value)  # VKRT_ERROR_INVALID_ARGUMENT, nothing launched
    assert len(tlas.trace_rays_static(rays[:0])) == 0
    want = tlas.trace_rays(rays[:1000])
    assert same_records(tlas.trace_rays_static(rays[:1000]), want)  # still usable after the refusal
    # device pointers for both buffers (the caller's own allocations; the entry point copies with cudaMemcpyDefault)
    import ctypes as C

    import torch

    drays = torch.from_numpy(rays[:1000].view(np.uint8).reshape(-1).copy()).cuda()
    dhits = torch.full((1000 * 24,), 0xEE, dtype=torch.uint8, device="cuda")
    torch.cuda.synchronize()
    before = torch.cuda.memory_allocated()
    ctx.check(ctx.lib.vkrt_trace_rays_static(ctx.handle, tlas.handle, C.c_void_p(drays.data_ptr()), 1000, 0, C.c_void_p(dhits.data_ptr())))
    assert torch.cuda.memory_allocated() == before
    got = dhits.cpu().numpy().view(B.HIT_DTYPE)
    assert same_records(got, want)
    del drays, dhits
    tlas.close(), blas.close(), ctx.close()
