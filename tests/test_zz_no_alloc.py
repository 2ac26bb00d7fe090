"""Allocation-free entry point `slider_guidance_loss_noalloc` (include/slider_b200.h): same results as
`slider_guidance_loss`, scratch in a file-scope `static __device__` array of the library, no allocation by the call.
No counterpart in the reference.  Collected last on purpose (file name).
Passes alone in a fresh process and after the other tests."""
import ctypes as C

import pytest
import torch

pytestmark = pytest.mark.gpu


def _call_pair(n, dtype, sign, guided, seed):
    """(loss, grad) through the existing entry point (caller workspace) and through the no-alloc one, same inputs."""
    from sliders_conceptmod_b200 import _lib
    h = _lib.lib()
    gen = torch.Generator(device="cuda").manual_seed(seed)
    t, p, nn, q = (torch.randn(n, device="cuda", generator=gen).to(dtype) for _ in range(4))
    code = _lib.dtype_code(dtype)
    st = _lib.stream_ptr()
    ws = torch.zeros(h.slider_loss_workspace_bytes(), dtype=torch.uint8, device="cuda")
    loss_a = torch.full((1,), -1.0, device="cuda")
    loss_b = torch.full((1,), -2.0, device="cuda")
    grad_a, grad_b = torch.empty_like(t), torch.empty_like(t)
    pp, qq = (p.data_ptr(), q.data_ptr()) if guided else (None, None)
    _lib.check(h.slider_guidance_loss(loss_a.data_ptr(), grad_a.data_ptr(), t.data_ptr(), pp, nn.data_ptr(), qq, n, 3.0,
                                      sign, code, ws.data_ptr(), ws.numel(), st))
    torch.cuda.synchronize()
    before = torch.cuda.memory_allocated()
    _lib.check(h.slider_guidance_loss_noalloc(loss_b.data_ptr(), grad_b.data_ptr(), t.data_ptr(), pp, nn.data_ptr(), qq,
                                              n, 3.0, sign, code, st))
    torch.cuda.synchronize()
    assert torch.cuda.memory_allocated() == before          # nothing went through torch's allocator either
    return loss_a, grad_a, loss_b, grad_b


@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
def test_noalloc_matches_existing_entry_point_twice_in_a_row(dtype):
    # two calls in a row share the built-in array (it keeps the first call's partial sums and ticket state);
    # different sizes / grids on purpose: 16 384 = SD-1.4 latents, 1 048 583 = many blocks + a ragged tail
    for k, (n, sign, guided) in enumerate([(1 << 20 | 7, +1, True), (16384, -1, True), (65536, +1, False)]):
        la, ga, lb, gb = _call_pair(n, dtype, sign, guided, seed=100 + k)
        # same kernel, same grid, same deterministic reduction order: bit-identical (the tests' tolerance for the
        # elementwise kernels is <= 2 fp32 ulp / 1 bf16 ulp; equality is within it)
        assert torch.equal(la, lb), (k, la.item(), lb.item())
        assert torch.equal(ga, gb), k
        assert torch.isfinite(lb).all() and lb.item() > 0


def test_noalloc_rejects_sizes_beyond_the_limit():
    from sliders_conceptmod_b200 import _lib
    h = _lib.lib()
    limit = h.slider_guidance_loss_noalloc_max_elems()
    assert limit == 1 << 28
    x = torch.zeros(16, device="cuda")
    out = torch.zeros(1, device="cuda")
    rc = h.slider_guidance_loss_noalloc(out.data_ptr(), None, x.data_ptr(), None, x.data_ptr(), None, limit + 1, 1.0, 1,
                                        _lib.dtype_code(torch.float32), _lib.stream_ptr())
    assert rc == 2 and b"limit" in h.slider_last_error() and str(limit).encode() in h.slider_last_error()
    rc = h.slider_guidance_loss_noalloc(out.data_ptr(), None, x.data_ptr(), None, x.data_ptr(), None, 0, 1.0, 1,
                                        _lib.dtype_code(torch.float32), _lib.stream_ptr())
    assert rc == 1
