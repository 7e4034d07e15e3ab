"""Host-side arithmetic that decides what the kernels are asked to do, checked on the CPU."""
import itertools

import pytest
import torch
import torch.nn.functional as F


@pytest.mark.parametrize('k,pad', [(4, 1), (3, 1), (4, 2), (5, 2), (2, 0), (3, 0), (6, 2)])
def test_stride2_data_gradient_as_four_parity_class_convolutions(k, pad):
    """functional._parity_taps: dx[2m + a] = sum_t dy[m + t - left] * w[k0 + 2 (T - 1 - t)] per axis.  Emulated with
    torch convolutions on the CPU (zero rows outside dy on every side, as TMA's out-of-bounds fill provides) and
    compared with torch's data gradient -- the kernel path (functional._dgrad_stride2) feeds exactly these taps,
    paddings and output sizes to imx_conv2d_igemm_fprop_window_bf16."""
    from imaginaire_b200.functional import _parity_taps
    g = torch.Generator().manual_seed(k * 10 + pad)
    for h, w in itertools.product((9, 10), (12, 13)):
        ho, wo = (h + 2 * pad - k) // 2 + 1, (w + 2 * pad - k) // 2 + 1
        if ho <= 0 or wo <= 0:
            continue
        cin, cout, n = 3, 5, 2
        wt = torch.randn(cout, cin, k, k, generator=g, dtype=torch.float64)
        dy = torch.randn(n, cout, ho, wo, generator=g, dtype=torch.float64)
        ref = torch.nn.grad.conv2d_input((n, cin, h, w), wt, dy, stride=2, padding=pad)
        dx = torch.full((n, cin, h, w), float('nan'), dtype=torch.float64)
        for a, b in itertools.product((0, 1), (0, 1)):
            ty, tx = _parity_taps(k, pad, a), _parity_taps(k, pad, b)
            assert ty is not None and tx is not None
            (ky0, nty, py), (kx0, ntx, px) = ty, tx
            ma, mb = (h - a + 1) // 2, (w - b + 1) // 2
            # correlation filter of the class: wsub[t] = w[k0 + 2 (T - 1 - t)], contraction over Cout
            sub = wt[:, :, ky0::2, kx0::2].flip(2, 3).permute(1, 0, 2, 3)          # [Cin, Cout, T, T]
            assert sub.shape[2:] == (nty, ntx)
            # out[m] = sum_t in[m + t - left]: pad left by `left`, right by whatever ma rows need
            need_h, need_w = ma + nty - 1, mb + ntx - 1
            padded = F.pad(dy, (px, max(0, need_w - px - wo), py, max(0, need_h - py - ho)))[:, :, :need_h, :need_w]
            dx[:, :, a::2, b::2] = F.conv2d(padded, sub)
        assert torch.allclose(dx, ref, atol=1e-10), (h, w)


def test_parity_taps_refuses_what_a_class_cannot_express():
    from imaginaire_b200.functional import _parity_taps
    assert _parity_taps(1, 0, 1) is None          # a 1x1 stride-2 filter never reaches odd pixels
    assert _parity_taps(3, 3, 0) is None          # padding beyond the filter: negative left padding
    assert _parity_taps(4, 1, 0) == (1, 2, 1) and _parity_taps(4, 1, 1) == (0, 2, 0)
    assert _parity_taps(3, 1, 0) == (1, 1, 0) and _parity_taps(3, 1, 1) == (0, 2, 0)
