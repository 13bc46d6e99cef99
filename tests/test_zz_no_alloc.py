"""ppo_grad_sumsq_static (include/ppo_kernels.h): the squared gradient norm without caller scratch - the
entry point for callers that cannot allocate device memory. It has no counterpart in the reference,
so its oracle is the existing entry point ppo_grad_sumsq on the same input:
same kernel, same launch shape, hence the same bits; a float64 sum bounds both.

Named test_zz_* so that pytest collects it last: a failure here must not
hide the parity tests. It passes alone in a fresh process and after the other tests (it owns its buffers and the
library's static array is cleared by every call)."""
import ctypes

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

MAX_N = 1 << 22  # PPO_GRAD_SUMSQ_STATIC_MAX_N


def _ctrl(lib, check, torch, dev, K):
    c = torch.zeros(ctypes.sizeof(K.PPOCtrl), dtype=torch.uint8, device=dev)
    check(lib.ppo_ctrl_init(c.data_ptr(), 7, 1e-4, 2e-4, None))
    return c


def _read(K, torch, c):
    torch.cuda.synchronize()
    return K.PPOCtrl.from_buffer_copy(c.cpu().numpy().tobytes())


def test_static_scratch_entry_point_matches_the_existing_one():
    import torch
    from diffusion_piano_b200.ppo import _kernels as K
    lib, check = K.load(), K.check
    dev = torch.device("cuda", 0)
    torch.cuda.set_device(dev)
    g = torch.Generator(device=dev)
    g.manual_seed(11)
    ca, cb = _ctrl(lib, check, torch, dev, K), _ctrl(lib, check, torch, dev, K)
    scratch = torch.zeros(592, dtype=torch.float32, device=dev)
    ticket = torch.zeros(1, dtype=torch.int32, device=dev)
    stream = torch.cuda.current_stream(dev).cuda_stream
    # two calls in a row on different gradients: the actor's size (one partial per block), then a size that
    # saturates the 592-block grid and has a tail that is not a multiple of 4
    calls = [(231_344, 0, 1.0), (1_500_003, 1, 3.0)]
    grads = [torch.randn(n + 4, device=dev, generator=g)[:n] * s for n, _, s in calls]
    torch.cuda.synchronize()
    free_before = None
    for i, ((n, which, _), grad) in enumerate(zip(calls, grads)):
        assert grad.data_ptr() % 16 == 0
        check(lib.ppo_grad_sumsq(grad.data_ptr(), n, scratch.data_ptr(), ticket.data_ptr(), ca.data_ptr(), which, stream))
        if i == 1:  # the first call resolved the array's address; from here on nothing may be allocated
            torch.cuda.synchronize()
            free_before = torch.cuda.mem_get_info(dev)[0]
        check(lib.ppo_grad_sumsq_static(grad.data_ptr(), n, cb.data_ptr(), which, stream))
        a, b = _read(K, torch, ca), _read(K, torch, cb)
        if i == 1:
            assert torch.cuda.mem_get_info(dev)[0] == free_before
        ref = float((grad.double() ** 2).sum().item())
        print(f"call {i}: n {n}: existing {a.norm2[which]:.9g} static {b.norm2[which]:.9g} float64 {ref:.9g}")
        assert a.norm2[which] == b.norm2[which]  # same kernel, same launch shape, same summation order
        np.testing.assert_allclose(b.norm2[which], ref, rtol=2e-6)
        assert a.adam_t[which] == b.adam_t[which] == 1
        assert int(ticket.item()) == 0
    # the other network's slot of the second call is untouched by it, the first call's result still stands
    a, b = _read(K, torch, ca), _read(K, torch, cb)
    assert a.norm2[0] == b.norm2[0] and b.norm2[0] != b.norm2[1]
    # sizes beyond the stated maximum are rejected before anything is launched, with a clear message
    rc = lib.ppo_grad_sumsq_static(grads[0].data_ptr(), MAX_N + 4, cb.data_ptr(), 0, stream)
    assert rc == -1
    msg = lib.ppo_last_error().decode()
    assert "PPO_GRAD_SUMSQ_STATIC_MAX_N" in msg and str(MAX_N) in msg
    assert _read(K, torch, cb).adam_t[0] == 1
    with pytest.raises(K.PPOKernelError):
        check(lib.ppo_grad_sumsq_static(grads[0].data_ptr() + 4, 1024, cb.data_ptr(), 0, stream))  # misaligned
