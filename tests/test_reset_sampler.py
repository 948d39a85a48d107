"""Distribution parity of the Philox reset sampler and random-walk command against the REFERENCE's generators.

The reference draws from three MT19937 streams per env (rate_env.py:64,73-77); a counter-based generator cannot replay them
(RNG row), so what is pinned here is what the draws are distributed like: tests/golden/reset_distributions.npz
holds, for 100 000 resets of the unmodified RateControlEnv per (difficulty, command type), exact supports (active-axis
sets, signs) and fine histograms (= the empirical CDF on a grid) of every drawn quantity, and for 400 000 calls of
RateCommandGenerator.generate_random_walk the normalised delta (tests/golden/make_golden.py::gen_reset_distributions).
The restatement's Philox sampler (which the CUDA reset kernel reproduces bit for bit, tests/test_gpu_env_f64.py::
test_reset_sampler_matches_oracle_bit_for_bit) is compared with them and with the closed forms of generators.py:47-214.

Statistics: two-sample Kolmogorov-Smirnov on the grid, D <= c * sqrt((n+m)/(n*m)) + 1/bins with c = 2.2 (alpha ~ 1e-4 per
test; the seeds are fixed, so the outcome is deterministic); chi-square for the active-axis sets; a 4.5-sigma binomial bound
for signs."""
import os

import numpy as np
import pytest

from conftest import GOLDEN
from oracle import oracle as orc

N = 400_000
SCALE = {"easy": 0.3, "medium": 0.5, "hard": 0.7}
CTYPE = {"step": orc.CMD_STEP, "ramp": orc.CMD_RAMP, "sine": orc.CMD_SINE}
MAX_RATES = np.radians([180.0, 180.0, 160.0])
C_ALPHA = 2.2


@pytest.fixture(scope="module")
def ref():
    return np.load(os.path.join(GOLDEN, "reset_distributions.npz"))


def ks_on_grid(ref_hist, ours, lo, hi):
    """sup over the grid's bin edges of |F_ref - F_ours|, and the bound it must stay under."""
    bins = len(ref_hist)
    h = np.histogram(ours, bins=bins, range=(lo, hi))[0]
    assert h.sum() == len(ours), "a draw left the reference's support"
    n, m = int(ref_hist.sum()), len(ours)
    d = float(np.max(np.abs(np.cumsum(ref_hist) / n - np.cumsum(h) / m)))
    return d, C_ALPHA * np.sqrt((n + m) / (n * m)) + 1.0 / bins


def uniform_ks(x, lo, hi):
    """one-sample KS of x against U(lo, hi)"""
    u = np.sort((x - lo) / (hi - lo))
    k = np.arange(1, len(u) + 1)
    return float(max(np.max(k / len(u) - u), np.max(u - (k - 1) / len(u)))), C_ALPHA / np.sqrt(len(u))


def draw(difficulty, ctype, seed):
    env = orc.Env(N, seed=seed, difficulty=SCALE[difficulty], command_type=CTYPE[ctype], id_offset=12345)
    env.reset()
    return env


def test_envelope_draws_match_the_reference_sampler(ref):
    """FlightEnvelopeSampler.sample (generators.py:190-214) through RateControlEnv.reset (rate_env.py:170-185)."""
    env = draw("medium", "step", seed=11)
    s = env.state
    q = np.stack([s[3], -s[2], s[6], s[7], s[8], s[9], s[10], s[11]])
    assert np.all(s[0] == 0) and np.all(s[1] == 0) and np.all(s[4] == 0) and np.all(s[5] == 0) and np.all(s[12] == 0)
    names = ["airspeed", "altitude", "roll", "pitch", "yaw", "p", "q", "r"]
    for j, name in enumerate(names):
        lo, hi = ref["env_support"][j]
        d, bound = ks_on_grid(ref["env_hist"][j], q[j], lo, hi)
        assert d <= bound, (name, d, bound)
        d1, b1 = uniform_ks(q[j], lo, hi)               # the closed form: uniform on the stated range
        assert d1 <= b1, (name, d1, b1)
    # the eight draws of one reset are independent in the reference (consecutive MT19937 outputs): so are the Philox words
    corr = np.corrcoef(q)
    assert np.max(np.abs(corr - np.eye(8))) <= 5.0 / np.sqrt(N)
    assert float(ref["env_corr_max_offdiag"]) <= 5.0 / np.sqrt(1e5)
    # airspeed / altitude derived outputs of the reset state (rate_env.py:181-182)
    assert np.allclose(env.derived[0], s[3]) and np.array_equal(env.derived[1], -s[2])


@pytest.mark.parametrize("difficulty", ["easy", "medium", "hard"])
@pytest.mark.parametrize("ctype", ["step", "ramp", "sine"])
def test_command_draws_match_the_reference_generator(ref, difficulty, ctype):
    """generate_step_command / generate_ramp_command / generate_sine_command (generators.py:47-132) as called by
    _generate_new_command (rate_env.py:302-331)."""
    key = f"{difficulty}/{ctype}"
    env = draw(difficulty, ctype, seed=100 + 7 * list(SCALE).index(difficulty) + CTYPE[ctype])
    if ctype == "step":
        cmd = env.rate_cmd.T
        assert np.all(env.sched == 0)
    else:
        cmd = env.sched[:3].T
        assert np.all(env.rate_cmd == 0)                                   # ramp start / sine at t = 0
        assert np.all(env.sched[4:] == 0)
    active = cmd != 0.0
    # exact support of the active-axis sets, and their frequencies (chi-square, 6 or 5 degrees of freedom)
    subset = np.bincount((active * np.array([1, 2, 4])).sum(axis=1), minlength=8)
    r_subset = ref[key + "/subset"]
    assert np.array_equal(subset > 0, r_subset > 0), (subset, r_subset)
    nz = r_subset > 0
    expect = r_subset[nz] / r_subset.sum() * N
    pooled = (subset[nz] + r_subset[nz]) / (N + r_subset.sum())            # two-sample chi-square on pooled proportions
    chi2 = float(np.sum((subset[nz] - N * pooled) ** 2 / (N * pooled) + (r_subset[nz] - r_subset.sum() * pooled) ** 2
                        / (r_subset.sum() * pooled)))
    assert chi2 <= 30.0, (chi2, subset, expect)                            # chi2(6) at 1e-4 is 27.9
    # closed form: number of axes uniform on {1,2,3} ({1,2} for sine), axes without replacement
    p_closed = {"step": [0, 1 / 9, 1 / 9, 1 / 9, 1 / 9, 1 / 9, 1 / 9, 1 / 3], "ramp": [0, 1 / 9, 1 / 9, 1 / 9, 1 / 9, 1 / 9, 1 / 9, 1 / 3],
                "sine": [0, 1 / 6, 1 / 6, 1 / 6, 1 / 6, 1 / 6, 1 / 6, 0]}[ctype]
    sig = np.sqrt(N * np.array(p_closed) * (1 - np.array(p_closed)))
    assert np.all(np.abs(subset - N * np.array(p_closed)) <= 4.5 * sig + 1e-9), subset
    # signs: +-1 with equal probability for step / ramp, always positive amplitudes for sine (generators.py:66,95,128)
    for a in range(3):
        n_act, n_pos = int(active[:, a].sum()), int((cmd[:, a] > 0).sum())
        r_act, r_pos = int(ref[key + "/active"][a]), int(ref[key + "/positive"][a])
        if ctype == "sine":
            assert n_pos == n_act and r_pos == r_act
        else:
            assert abs(n_pos - 0.5 * n_act) <= 4.5 * np.sqrt(0.25 * n_act)
            assert abs(n_pos / n_act - r_pos / r_act) <= 4.5 * np.sqrt(0.25 / n_act + 0.25 / r_act)
    # magnitudes: U(0.3, 1) * [0.5 *] difficulty_scale * max_rate per active axis
    norm = np.abs(cmd) / ((0.5 if ctype == "sine" else 1.0) * SCALE[difficulty] * MAX_RATES)
    for a in range(3):
        x = norm[active[:, a], a]
        d, bound = ks_on_grid(ref[key + "/mag_hist"][a], x, 0.3, 1.0)
        assert d <= bound, (a, d, bound)
        d1, b1 = uniform_ks(x, 0.3, 1.0)
        assert d1 <= b1, (a, d1, b1)
        assert abs(x.mean() - float(ref[key + "/mag_mean"][a])) <= 5 * 0.7 / np.sqrt(12) * np.sqrt(1 / len(x) + 1 / int(ref[key + "/active"][a]))
    if ctype == "sine":
        d, bound = ks_on_grid(ref[key + "/freq_hist"], env.sched[3], 0.1, 2.0)
        assert d <= bound, (d, bound)
        d1, b1 = uniform_ks(env.sched[3], 0.1, 2.0)
        assert d1 <= b1
    else:
        assert np.all(env.sched[3] == 0)


@pytest.mark.parametrize("difficulty", ["easy", "medium", "hard"])
def test_random_walk_normals_match_the_reference_generator(ref, difficulty):
    """generate_random_walk (generators.py:145-164): delta = randn(3) * 0.1 * sqrt(dt) * scale * max_rates.  The Box-Muller
    normals of the aircraft's Philox stream against the reference's MT19937 randn, and the per-axis delta spread."""
    n_air, n_steps = 2000, 100
    z = np.stack([orc.walk_normals(21, gid, 1 + gid % 3, t) for gid in range(n_air) for t in range(n_steps)])
    key = f"{difficulty}/walk"
    assert int(ref[key + "/z_outside"]) == 0 and np.all(np.abs(z) < 6.0)
    for a in range(3):
        d, bound = ks_on_grid(ref[key + "/z_hist"][a], z[:, a], -6.0, 6.0)
        assert d <= bound, (a, d, bound)
    m = len(z)
    assert np.all(np.abs(z.mean(axis=0)) <= 4.5 / np.sqrt(m)) and np.all(np.abs(z.std(axis=0) - 1) <= 4.5 / np.sqrt(2 * m))
    assert np.max(np.abs(np.corrcoef(z.T) - np.eye(3))) <= 5 / np.sqrt(m)
    assert np.max(np.abs(ref[key + "/z_corr"] - np.eye(3))) <= 5 / np.sqrt(4e5)
    # consecutive steps of one aircraft are independent
    zz = z.reshape(n_air, n_steps, 3)
    lag = np.mean(zz[:, 1:, :] * zz[:, :-1, :])
    assert abs(lag) <= 5 / np.sqrt(zz[:, 1:, :].size)
    # the reference's delta spread is 0.1 * sqrt(0.02) * scale * max_rate; the env's walk (oracle, Philox) has the same
    closed = 0.1 * np.sqrt(0.02) * SCALE[difficulty] * MAX_RATES
    assert np.all(np.abs(ref[key + "/delta_std"] / closed - 1) <= 4.5 / np.sqrt(2 * 4e5))
    env = orc.Env(50_000, seed=5, difficulty=SCALE[difficulty], command_type=orc.CMD_RANDOM, controller=orc.CTRL_RATE_PID,
                  auto_reset=False)
    env.reset()
    assert np.all(env.rate_cmd == 0)                                        # rate_env.py:333-336
    env.step()
    d1 = env.rate_cmd.copy()                                                # first delta: no clip can bind yet
    assert np.all(np.abs(d1.std(axis=1) / closed - 1) <= 4.5 / np.sqrt(2 * 50_000))
    for a in range(3):
        dd, bound = ks_on_grid(ref[key + "/z_hist"][a], d1[a] / closed[a], -6.0, 6.0)
        assert dd <= bound
