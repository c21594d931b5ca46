"""-m gpu: dglb200_gather_rows_convert, the converting row gather of the reduced-precision wire mode
of the halo exchange (dgl_b200/dist.py, DGL_B200_WIRE=bf16).  Integer-exact expectations:
float32 -> bfloat16 is round-to-nearest-even (== torch's cast), bfloat16 -> float32 is exact."""
import pytest
import torch

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("feat", [128, 8, 5, 13, 64 + 3])
@pytest.mark.parametrize("with_idx", [True, False])
def test_gather_rows_convert_bit_exact(feat, with_idx):
    from dgl_b200 import ops
    gen = torch.Generator(device="cuda").manual_seed(feat)
    n = 1000
    x = torch.randn(n, feat, device="cuda", generator=gen) * 37.0
    idx = torch.randint(0, n, (700,), device="cuda", generator=gen) if with_idx else None
    src = x[idx] if with_idx else x
    out = torch.full((src.shape[0], feat), float("nan"), dtype=torch.bfloat16, device="cuda")
    ops.gather_rows_convert(x, idx, out)
    assert torch.equal(out.view(torch.int16), src.to(torch.bfloat16).view(torch.int16))
    back = torch.full((src.shape[0], feat), float("nan"), device="cuda")
    ops.gather_rows_convert(out, None, back)
    assert torch.equal(back, out.float())
    # widening with a gather, out of a column block of a wider matrix, into a row slice
    wide = torch.zeros(n, feat + 8, dtype=torch.bfloat16, device="cuda")
    wide[:, :feat] = x.to(torch.bfloat16)
    pick = torch.randint(0, n, (50,), device="cuda", generator=gen)
    big = torch.zeros(80, feat, device="cuda")
    ops.gather_rows_convert(wide[:, :feat], pick, big[10:60])
    assert torch.equal(big[10:60], wide[pick, :feat].float())
    assert float(big[:10].abs().sum()) == 0.0 and float(big[60:].abs().sum()) == 0.0


def test_gather_rows_convert_rejects_bad_arguments():
    from dgl_b200 import ops
    from dgl_b200.base import DGLError
    x = torch.randn(4, 8, device="cuda")
    with pytest.raises(DGLError):
        ops.gather_rows_convert(x, None, torch.empty(4, 8, device="cuda"))              # same dtype
    with pytest.raises(DGLError):
        ops.gather_rows_convert(x, None, torch.empty(4, 7, dtype=torch.bfloat16, device="cuda"))
    with pytest.raises(DGLError):
        ops.gather_rows_convert(x, None, torch.empty(3, 8, dtype=torch.bfloat16, device="cuda"))
