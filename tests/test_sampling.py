"""Top-k / top-p sampling (Generate.cpp:45-143).  The reference seeds its generator from std::random_device on every call, so what
can be pinned is the DISTRIBUTION: the restatement's candidate probabilities against the empirical frequencies of the compiled
reference classes (CPU), and the device kernel's candidates / probabilities / draws against the restatement (GPU)."""
import numpy as np
import pytest

from oracle import pyoracle as po


def _row(n, seed, scale=3.0):
    return (np.random.default_rng(seed).standard_normal(n) * scale).astype(np.float32)


def _softmax(x):
    e = np.exp(x.astype(np.float64) - x.max())
    return (e / e.sum()).astype(np.float32)


@pytest.mark.parametrize("mode,arg,T", [(0, 5, 0.8), (0, 12, 1.5), (1, 0.6, 1.0), (1, 0.9, 0.05)])
def test_restatement_matches_reference_sampler_frequencies(mode, arg, T):
    if po.cpuref() is None:
        pytest.skip("oracle/_ref/libmllm_cpu_ref.so not built (needs /root/reference)")
    x = _row(300, 11 + int(arg * 10))
    if mode == 1:
        x = _softmax(x)
    draws = 40000
    counts = po.ref_sample_hist(mode, x, arg, T, draws)
    ids, p = po.sample_topk_candidates(x, arg, T) if mode == 0 else po.sample_topp_candidates(x, arg, T)
    assert set(np.nonzero(counts)[0]) <= set(ids.tolist())           # the reference never leaves the candidate set
    freq = counts[ids] / draws
    sigma = np.sqrt(np.maximum(p * (1 - p), 1e-9) / draws)
    assert np.all(np.abs(freq - p) <= 5 * sigma + 1e-4), (freq, p)
    assert abs(float(p.astype(np.float64).sum()) - 1.0) < 1e-5


def test_draw_is_the_inverse_cdf():
    p = np.array([0.5, 0.25, 0.125, 0.125], dtype=np.float32)
    assert [po.sample_draw(p, u) for u in (0.0, 0.49, 0.5, 0.74, 0.75, 0.874, 0.876, 0.9999)] == [0, 0, 1, 1, 2, 2, 3, 3]


@pytest.mark.gpu
@pytest.mark.parametrize("n,mode,arg,T", [(32000, 0, 5, 0.8), (151936, 0, 40, 1.0), (1000, 0, 2, 0.3), (32000, 1, 0.8, 1.0), (32000, 1, 0.95, 0.5), (4096, 1, 0.3, 1.0)])
def test_device_sampler_against_restatement(cu, n, mode, arg, T):
    import torch
    rows = 3
    X = np.stack([_row(n, 5 * r + n % 7, scale=4.0) for r in range(rows)])
    if mode == 1:
        X = np.stack([_softmax(x) for x in X])
    X[1, 17] = X[1, 900]  # a tie: lowest id first
    u = np.array([0.03, 0.5, 0.97], dtype=np.float32)
    ids, cp, ci, nc = cu.sample(torch.from_numpy(X).cuda(), torch.from_numpy(u).cuda(), k=arg if mode == 0 else None, p=arg if mode == 1 else None,
                                temperature=T, cand_cap=1024)
    ids, cp, ci, nc = ids.cpu().numpy(), cp.cpu().numpy(), ci.cpu().numpy(), nc.cpu().numpy()
    for r in range(rows):
        wi, wp = po.sample_topk_candidates(X[r], arg, T) if mode == 0 else po.sample_topp_candidates(X[r], arg, T)
        assert nc[r] == len(wi) and np.array_equal(ci[r, :nc[r]], wi)
        assert np.array_equal(cp[r, :nc[r]].view(np.uint32), wp.view(np.uint32)), np.abs(cp[r, :nc[r]] - wp).max()   # bit-exact probabilities
        assert int(ids[r]) == int(wi[po.sample_draw(wp, float(u[r]))])
