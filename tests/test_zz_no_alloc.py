"""vnb_gn_stats_static: the allocation-free variant of vnb_gn_stats (its chunk-wise workspace is a file-scope static
__device__ array inside the library; no counterpart in the reference).
Collected last on purpose.  Called through the C ABI, compared with vnb_gn_stats on two calls
in a row; passes alone in a fresh process and after the other tests (nothing in the shared array needs clearing)."""
import pytest
import torch

from conftest import rel_err

pytestmark = pytest.mark.gpu


def _frames(B, C, T, seed):
    g = torch.Generator().manual_seed(seed)
    # channels-last rows (B, T, C), the layout every kernel of the path reads; offset mean so that rstd matters
    return (torch.randn(B, T, C, generator=g) * 1.7 + 0.3).cuda()


def _both(h, B, C, T, groups, eps=1e-5):
    from vae_npvc_b200 import _C
    ws = torch.empty(_C.call('gn_stats_ws_floats', B, groups), dtype=torch.float32, device='cuda')
    mean0, rstd0 = torch.empty(B, groups, device='cuda'), torch.empty(B, groups, device='cuda')
    mean1, rstd1 = torch.full((B, groups), float('nan'), device='cuda'), torch.full((B, groups), float('nan'), device='cuda')
    _C.call('gn_stats', h, B, C, T, groups, eps, ws, mean0, rstd0)
    torch.cuda.synchronize()
    before = torch.cuda.memory_allocated()
    _C.call('gn_stats_static', h, B, C, T, groups, eps, mean1, rstd1)        # outputs were allocated by the caller, above
    torch.cuda.synchronize()
    assert torch.cuda.memory_allocated() == before
    return (mean0, rstd0), (mean1, rstd1)


def test_gn_stats_static_matches_gn_stats_on_two_calls_in_a_row():
    from vae_npvc_b200 import _C
    _C.require_cuda()
    eps = 1e-5
    # first call: the VCC2020 decoder shape (GroupNorm(2, 1024) over 16 x 128 frames); second: a smaller, ragged shape
    # that reuses the front of the same static array while the rest still holds the first call's partial sums
    for (B, C, T, groups, seed) in [(16, 1024, 128, 2, 1), (5, 96, 77, 1, 2), (16, 1024, 128, 2, 3)]:
        h = _frames(B, C, T, seed)
        (m0, r0), (m1, r1) = _both(h, B, C, T, groups, eps)
        # same kernels, same workspace layout; the chunk kernel sums with shared-memory float atomics, whose order is
        # not fixed, so two runs of EITHER entry point agree to fp32 rounding, not bit for bit
        assert rel_err(m1, m0) < 1e-6 and rel_err(r1, r0) < 1e-6
        x = h.double().view(B, T, groups, C // groups)
        mean = x.mean(dim=(1, 3))
        var = x.var(dim=(1, 3), unbiased=False)
        assert rel_err(m1, mean) < 1e-5
        assert rel_err(r1, (var + eps).rsqrt()) < 1e-5


def test_gn_stats_static_rejects_shapes_beyond_the_static_workspace():
    from vae_npvc_b200 import _C
    lib = _C.load()
    assert lib.vnb_gn_stats_static_max_floats() == 65536
    B, C, T, groups = 513, 8, 4, 1                                            # B * groups = 513 > 512
    h = _frames(B, C, T, 4)
    mean, rstd = torch.empty(B, groups, device='cuda'), torch.empty(B, groups, device='cuda')
    with pytest.raises(RuntimeError, match='static workspace holds 65536'):
        _C.call('gn_stats_static', h, B, C, T, groups, 1e-5, mean, rstd)
    # the largest shape it takes
    B = 512
    h = _frames(B, C, T, 5)
    (m0, r0), (m1, r1) = _both(h, B, C, T, groups)
    assert rel_err(m1, m0) < 1e-6 and rel_err(r1, r0) < 1e-6
