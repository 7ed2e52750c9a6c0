"""GPU checks of the bf16 tcgen05 path against torch fp32 references of the same op evaluated on the
SAME bf16-rounded operands (so the only difference is fp32 accumulation order): rtol 2e-3."""
import numpy as np
import pytest
import torch
import torch.nn.functional as F

pytestmark = pytest.mark.gpu


def _bf(x):
    return x.to(torch.bfloat16).to(torch.float32)


@pytest.mark.parametrize("B,Cin,M,T,K,dil", [(2, 192, 384, 430, 5, 1), (3, 80, 192, 77, 1, 1), (1, 192, 160, 128, 1, 1),
                                              (2, 192, 192, 300, 3, 2), (2, 384, 192, 129, 5, 1), (1, 64, 16, 5, 5, 3)])
def test_tc_conv1d_matches_torch(B, Cin, M, T, K, dil):
    from glowtts_b200 import ops
    gen = torch.Generator(device="cuda").manual_seed(B * 1000 + T)
    x = torch.randn(B, Cin, T, device="cuda", generator=gen)
    w = torch.randn(M, Cin, K, device="cuda", generator=gen) / (Cin * K) ** 0.5
    bias = torch.randn(M, device="cuda", generator=gen)
    lengths = torch.randint(T // 2, T + 1, (B,), device="cuda", generator=gen)
    mask = (torch.arange(T, device="cuda")[None, :] < lengths[:, None]).float()
    frames = ops.tc_frames(x)
    assert frames.shape == (B, T, (Cin + 63) // 64 * 64)
    np.testing.assert_array_equal(frames[:, :, :Cin].float().cpu().numpy(), _bf(x).permute(0, 2, 1).cpu().numpy())
    assert frames[:, :, Cin:].abs().sum().item() == 0
    wp = ops.tc_pack_weight(w)
    out32, out16 = ops.tc_conv1d(frames, wp, bias, dil, mask, want_f32=True, want_bf16=True)
    ref = F.conv1d(_bf(x), _bf(w), bias, padding=(K - 1) * dil // 2, dilation=dil) * mask[:, None, :]
    ref = ref.permute(0, 2, 1)
    torch.cuda.synchronize()
    np.testing.assert_allclose(out32.cpu().numpy(), ref.cpu().numpy(), rtol=2e-3, atol=2e-3)
    np.testing.assert_allclose(out16[:, :, :M].float().cpu().numpy(), ref.cpu().numpy(), rtol=1e-2, atol=1e-2)


def test_tc_conv1d_data_gradient_form():
    """conv^T: the packed transposed+flipped weights give the gradient w.r.t. the conv input."""
    from glowtts_b200 import ops
    gen = torch.Generator(device="cuda").manual_seed(7)
    B, Cin, M, T, K, dil = 2, 192, 384, 200, 5, 1
    x = torch.randn(B, Cin, T, device="cuda", generator=gen, requires_grad=True)
    w = torch.randn(M, Cin, K, device="cuda", generator=gen) / (Cin * K) ** 0.5
    dout = torch.randn(B, M, T, device="cuda", generator=gen)
    F.conv1d(x, _bf(w), None, padding=(K - 1) * dil // 2, dilation=dil).backward(_bf(dout))
    wpt = ops.tc_pack_weight(w, transpose=True)
    dx, _ = ops.tc_conv1d(ops.tc_frames(dout), wpt, None, dil)
    np.testing.assert_allclose(dx.cpu().numpy(), x.grad.permute(0, 2, 1).cpu().numpy(), rtol=2e-3, atol=2e-3)
