"""The weight-gradient entry point for callers that cannot allocate device memory.

Some users of this library run where a cudaMalloc is off limits at the moment the gradient
is needed - a server that hands out its whole device-memory budget at start-up and refuses
allocations afterwards is the typical case - so they cannot supply the split-K workspace of
imx_conv2d_igemm_wgrad_bf16.  imx_conv2d_igemm_wgrad_bf16_noalloc keeps that scratch in an
array inside the library instead.  This test calls it through the C ABI, twice in a row
with different layers (the second call finds the first call's leftovers in the scratch), and
checks each result against the ordinary entry point; it also checks that a layer too big for
the built-in array is refused with a clear message.

Collected last on purpose (file name): a -x run has already executed every other test.
"""
import pytest
import torch

from tests import helpers as H

pytestmark = pytest.mark.gpu
DEV = 'cuda:0'


def _nhwc(t):
    return t.to(DEV).bfloat16().contiguous(memory_format=torch.channels_last)


def _case(seed, n, cin, cout, h, w, k, stride, pad):
    g = torch.Generator().manual_seed(seed)
    ho, wo = (h + 2 * pad - k) // stride + 1, (w + 2 * pad - k) // stride + 1
    x = _nhwc(torch.randn(n, cin, h, w, generator=g))
    dy = _nhwc(torch.randn(n, cout, ho, wo, generator=g))
    return x, dy


def test_noalloc_wgrad_matches_workspace_entry_point_on_two_calls_in_a_row():
    from imaginaire_b200 import _C, ops
    lib = _C.lib()
    stream = torch.cuda.current_stream().cuda_stream
    # two different layers back to back: split-K with several splits, then a multi-tile layer
    cases = [(1, 2, 64, 512, 32, 64, 5, 1, 2), (2, 4, 256, 256, 32, 64, 3, 1, 1)]
    for seed, n, cin, cout, h, w, k, stride, pad in cases:
        x, dy = _case(seed, n, cin, cout, h, w, k, stride, pad)
        ref = ops.conv2d_wgrad(x, dy, cin, cout, k, k, stride, pad)           # existing entry point
        out = torch.full((cout, cin, k, k), float('nan'), device=DEV)        # caller-owned output
        # straight through the C ABI: pointers, sizes, stream - and no workspace argument
        lib.call('imx_conv2d_igemm_wgrad_bf16_noalloc', x.data_ptr(), dy.data_ptr(), out.data_ptr(),
                 n, h, w, cin, cin, cout, cout, k, k, stride, pad, stream)
        torch.cuda.synchronize()
        assert torch.isfinite(out).all()
        assert H.rel_err(out.cpu(), ref.cpu()) < 1e-4      # the tolerance of the wgrad parity tests
        # and against the CPU computation itself (fp32 accumulation of the exact bf16 products), so that the
        # numbers are checked and not only the scratch plumbing
        cpu = torch.nn.grad.conv2d_weight(x.float().cpu(), (cout, cin, k, k), dy.float().cpu(), stride=stride, padding=pad)
        assert H.rel_err(out.cpu(), cpu) < 1e-4
    # the tensor-level wrapper goes through the same entry point
    x, dy = _case(3, 1, 40, 72, 13, 11, 5, 1, 2)
    pad_c = lambda t, c: torch.nn.functional.pad(t.float(), (0, 0, 0, 0, 0, c - t.shape[1])).bfloat16() \
        .contiguous(memory_format=torch.channels_last)
    xp, dp = pad_c(x, 40), pad_c(dy, 72)
    assert H.rel_err(ops.conv2d_wgrad_noalloc(xp, dp, 40, 72, 5, 5, 1, 2).cpu(),
                     ops.conv2d_wgrad(xp, dp, 40, 72, 5, 5, 1, 2).cpu()) < 1e-4


def test_noalloc_wgrad_rejects_layers_beyond_the_builtin_scratch():
    from imaginaire_b200 import _C
    lib = _C.lib()
    # 3x3 2048 -> 2048: 9 * 2048 * 2048 fp32 partials = 151 MB > the 64 MiB built-in array
    need = lib.raw('imx_conv2d_wgrad_workspace_bytes')(8, 16, 32, 2048, 2048, 2048, 2048, 3, 3, 1, 1)
    assert need > 64 * 1024 * 1024
    dummy = torch.zeros(16, device=DEV)
    with pytest.raises(_C.ImxError, match='built-in') as err:
        lib.call('imx_conv2d_igemm_wgrad_bf16_noalloc', dummy.data_ptr(), dummy.data_ptr(), dummy.data_ptr(),
                 8, 16, 32, 2048, 2048, 2048, 2048, 3, 3, 1, 1, torch.cuda.current_stream().cuda_stream)
    assert 'code 10004' in str(err.value) and str(need) in str(err.value)
