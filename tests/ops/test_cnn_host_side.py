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
