"""Host-side fp32 -> bf16 staging (csrc/host_stage.cu): CPU code of the native library, testable without a GPU."""
import pytest
import torch

from elephas_b200.ops import build, native

pytestmark = pytest.mark.skipif(not native.available() and build.shutil.which('nvcc') is None,
                                reason='libelephas_b200.so not built and nvcc missing')


@pytest.mark.parametrize('rows,cols,ld,threads', [(1, 1, 8, 1), (37, 93, 96, 3), (1000, 784, 784, 4), (130, 4096, 4104, 8)])
def test_host_cast_is_bit_identical_to_torch(rows, cols, ld, threads):
    torch.manual_seed(rows)
    x = torch.randn(rows, cols) * 3
    x.view(-1)[::17] *= 1e-30        # denormal-range products
    x[0, 0] = float('inf')
    if cols > 2:
        x[0, 1], x[0, 2] = float('-inf'), float('nan')
    dst = torch.full((rows, ld), 5.0, dtype=torch.bfloat16)
    native.host_f32_to_bf16(x, dst, threads=threads)
    ref = x.to(torch.bfloat16)
    ok = ~torch.isnan(ref)
    assert torch.equal(dst[:, :cols].view(torch.int16)[ok], ref.view(torch.int16)[ok])
    assert torch.equal(torch.isnan(dst[:, :cols]), torch.isnan(ref))
    assert bool((dst[:, cols:] == 0).all())


def test_host_cast_of_a_row_slice_and_bad_arguments():
    x = torch.randn(64, 32)
    dst = torch.zeros(16, 32, dtype=torch.bfloat16)
    native.host_f32_to_bf16(x[8:24], dst, threads=2)
    assert torch.equal(dst, x[8:24].to(torch.bfloat16))
    with pytest.raises(native.NativeError):
        native.host_f32_to_bf16(x, torch.zeros(64, 32), threads=1)            # 4-byte destination
    with pytest.raises(native.NativeError):
        native.host_f32_to_bf16(x, torch.zeros(64, 16, dtype=torch.bfloat16))  # destination rows too short
