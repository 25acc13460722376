"""CPU: the C restatement (oracle-B) against the fixtures generated from the reference's own
Python (oracle/make_golden.py), plus unit checks of its pieces."""
import numpy as np
import pytest

import cf_oracle as co
import make_golden as mg
from helpers import GOLDEN_NAMES, load_fixture


def test_fixtures_present():
    assert len(GOLDEN_NAMES) >= 8


@pytest.mark.parametrize('name', GOLDEN_NAMES)
def test_oracle_matches_reference_fixture(name):
    fix = dict(np.load(f'{mg.GOLDEN_DIR}/{name}.npz'))
    dev = mg.replay_oracle(fix)
    assert dev['flags'] == 0                     # done / cost / truncated identical
    assert dev['obs'] < 1e-9 and dev['state'] < 1e-8 and dev['rew'] < 1e-9, dev


def test_grid_index_known_answers():
    # probed on the reference's own Grid.get_index
    idx = co.hj_index(np.array([[0.0] * 6, [2, -2, 3.1, 20, -20, 0.2]]))
    assert idx.tolist() == [[7, 7, 7, 7, 7, 7], [14, 0, 14, 14, 0, 7]]


def test_boltzmann_cdf():
    cdf = co.boltzmann_cdf()
    w = np.exp(-np.arange(0.0, 2.1, 0.1))
    ref = (w / np.sum(w)).cumsum()
    ref /= ref[-1]
    assert np.array_equal(cdf, ref)


def test_gae_matches_lfilter():
    import scipy.signal
    rng = np.random.default_rng(0)
    T, N, g, lam = 64, 5, 0.99, 0.95
    rew = rng.standard_normal((T, N)).astype(np.float32)
    val = rng.standard_normal((T, N)).astype(np.float32)
    done = np.zeros((T, N), np.uint8)
    last = rng.standard_normal(N).astype(np.float32)
    adv, tgt, ret = co.gae(rew, val, done, None, last, g, lam)

    def cumsum(x, d):   # algs/core.py:105-119
        return scipy.signal.lfilter([1], [1, float(-d)], x[::-1], axis=0)[::-1]
    for n in range(N):
        r = np.append(rew[:, n], last[n]).astype(np.float64)
        v = np.append(val[:, n], last[n]).astype(np.float64)
        deltas = r[:-1] + g * v[1:] - v[:-1]
        np.testing.assert_allclose(adv[:, n], cumsum(deltas, g * lam), rtol=2e-6, atol=2e-6)
        np.testing.assert_allclose(ret[:, n], cumsum(r, g)[:-1], rtol=2e-6, atol=2e-6)
        np.testing.assert_allclose(tgt[:, n], adv[:, n] + val[:, n], rtol=2e-6, atol=2e-6)
