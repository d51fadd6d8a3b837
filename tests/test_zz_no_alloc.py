"""mllmcu_linear_static_ws: the allocation-free Linear (scratch = a static __device__ array inside libmllmcu.so).

Collected last on purpose (file name).  Checked through the C ABI against the existing
entry point mllmcu_linear (caller-supplied workspace): the two run the same kernels, so the outputs must be bit-equal
(the tolerance the Linear tests use against the oracle, 2e-5 relative, is asserted as well and is implied by equality).
Two calls in a row share the static array; the second must not see anything of the first."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _weights(torch, n, k, seed):
    g = torch.Generator(device="cuda").manual_seed(seed)
    w = torch.randint(0, 256, (n, k // 32, 18), dtype=torch.uint8, device="cuda", generator=g)
    d = (torch.rand((n, k // 32), device="cuda", generator=g) * 0.01 + 0.001).to(torch.float16)
    w[..., 0:2] = d.view(torch.uint8).reshape(n, k // 32, 2)
    return w.reshape(n, -1).contiguous()


@pytest.mark.parametrize("shapes", [[(1, 512, 1024), (64, 256, 2048)], [(192, 384, 4096), (3, 128, 4096)]])
def test_static_ws_linear_matches_linear_twice_in_a_row(cu, shapes):
    import torch
    outs, wants = [], []
    prepared = []
    for i, (M, N, K) in enumerate(shapes):
        g = torch.Generator(device="cuda").manual_seed(100 + i)
        x = torch.randn((M, K), device="cuda", generator=g)
        W = _weights(torch, N, K, 200 + i)
        b = torch.randn((N,), device="cuda", generator=g) * 0.1
        prepared.append((W, x, b))
    assert cu.lib().mllmcu_linear_static_ws_capacity() == 16 << 20
    # the two static-scratch calls back to back on one stream, nothing in between
    for W, x, b in prepared:
        rc, y = cu.linear_static_ws(W, x, bias=b)
        assert rc == 0
        outs.append(y)
    torch.cuda.synchronize()
    for W, x, b in prepared:
        wants.append(cu.linear(W, x, bias=b))
    torch.cuda.synchronize()
    for y, want in zip(outs, wants):
        y, want = y.cpu().numpy(), want.cpu().numpy()
        assert np.isfinite(want).all()
        assert np.abs(y - want).max() <= 2e-5 * np.abs(want).max()
        assert np.array_equal(y, want)


def test_static_ws_linear_rejects_oversize_and_bad_args(cu):
    import torch
    cap = cu.lib().mllmcu_linear_static_ws_capacity()
    M, K = 4096, 8192
    assert cu.lib().mllmcu_linear_workspace_bytes(M, K) > cap
    x = torch.zeros((M, K), device="cuda")
    W = torch.zeros((32, K // 32 * 18), dtype=torch.uint8, device="cuda")
    rc, _ = cu.linear_static_ws(W, x)
    assert rc == -3                      # MLLMCU_EWS: beyond the static array
    rc, _ = cu.linear_static_ws(W[:, :18 * 4], x[:4, :128].contiguous(), wtype=5)
    assert rc == -2                      # unsupported weight type, same code as mllmcu_linear
    torch.cuda.synchronize()
