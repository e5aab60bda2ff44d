"""Host-side pieces of the native CNN path that need no GPU."""
import torch

from elephas_b200.ops import cnn


def test_flat_bf16_weight_copy_is_only_trusted_inside_a_training_step():
    """``weights_bf16`` hands out the engine's flat bf16 view only while ``fresh_weight_copy`` is active (the engine has
    just refreshed it); anywhere else - predict, evaluate, a step of another engine - it casts the fp32 values."""
    p = torch.nn.Parameter(torch.tensor([[1.0, 2.0], [3.0, 4.0]]))
    stale = torch.zeros(2, 2, dtype=torch.bfloat16)
    p._eb_w16 = stale
    assert torch.equal(cnn.weights_bf16(p).float(), p.detach())            # outside of a step: a fresh cast
    with cnn.fresh_weight_copy():
        assert cnn.weights_bf16(p) is stale                                  # inside: the view, no cast kernel
        with cnn.fresh_weight_copy():
            pass
        assert cnn.weights_bf16(p) is stale                                  # nesting restores the outer state
    assert torch.equal(cnn.weights_bf16(p).float(), p.detach())
    q = torch.nn.Parameter(torch.ones(3))
    with cnn.fresh_weight_copy():
        assert cnn.weights_bf16(q).dtype == torch.bfloat16                   # no view registered: cast


def test_pool_geometry_guard():
    x = torch.zeros(2, 9, 7, 16)
    assert cnn.max_pool_supported(x, (3, 3), (2, 2), (1, 1, 1, 1))
    assert not cnn.max_pool_supported(torch.zeros(2, 9, 7, 12), (3, 3), (2, 2))     # channels not a multiple of 8
    assert not cnn.max_pool_supported(x, (3, 3), (2, 1))                            # anisotropic stride
    assert not cnn.max_pool_supported(x, (2, 2), (2, 2), (2, 0, 0, 0))              # padding as large as the window


def test_conv_tile_patch_respects_the_tma_limits_for_arbitrary_geometries():
    """Property: the output-pixel patch of a convolution tile never exceeds the tile (128 rows), never asks a TMA box for
    more than 256 input elements along an edge, never exceeds the tensor, and wastes at most half of the tile rows unless
    the whole output is smaller than a tile."""
    from hypothesis import given, settings
    from hypothesis import strategies as st
    from elephas_b200.ops import native

    @settings(max_examples=200, deadline=None, derandomize=True)
    @given(st.integers(1, 300), st.integers(1, 230), st.integers(1, 230), st.sampled_from([1, 2, 3, 4]))
    def check(n, p, q, stride):
        pn, ph, pw = native.conv_patch(n, p, q, 128, stride)
        assert 1 <= pn <= n and 1 <= ph <= p and 1 <= pw <= q
        assert pn * ph * pw <= 128 and pn <= 256
        assert (ph - 1) * stride + 1 <= 256 + stride and (pw - 1) * stride + 1 <= 256 + stride
        tiles = -(-n // pn) * -(-p // ph) * -(-q // pw)
        if n * p * q >= 128:
            assert n * p * q / (tiles * 128.0) > 0.4

    check()
