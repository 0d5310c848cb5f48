"""oracle/philox.py, the CPU restatement of the in-kernel sampler noise: Random123 known-answer vectors for
Philox4x32-10 and the statistics the sampler needs (universe.py:39-41: unit normal, independent draws)."""
import numpy as np

from oracle import philox


def test_philox4x32_10_known_answers():
    # Random123 kat_vectors (philox4x32 10 rounds): counter, key -> output
    kat = [
        ((0, 0, 0, 0), (0, 0), (0x6627E8D5, 0xE169C58D, 0xBC57AC4C, 0x9B00DBD8)),
        ((0xFFFFFFFF,) * 4, (0xFFFFFFFF, 0xFFFFFFFF), (0x408F276D, 0x41C83B0E, 0xA20BC7C6, 0x6D5451FD)),
        ((0x243F6A88, 0x85A308D3, 0x13198A2E, 0x03707344), (0xA4093822, 0x299F31D0), (0xD16CFE09, 0x94FDCCEB, 0x5001E420, 0x24126EA1)),
    ]
    for ctr, key, want in kat:
        got = philox.philox4x32_10(*[[c] for c in ctr], *key)
        assert tuple(int(g[0]) for g in got) == want


def test_unit_normal_statistics_and_independence():
    z = philox.noise(seed=20240531, n_steps=3, B=3, Lp=65536)
    assert z.dtype == np.float32 and z.shape == (3, 3, 65536) and np.isfinite(z).all()
    assert abs(z.mean()) < 5e-3 and abs(z.std() - 1.0) < 5e-3
    assert abs((z ** 3).mean()) < 3e-2 and abs((z ** 4).mean() - 3.0) < 6e-2     # skewness, kurtosis
    flat = z.reshape(9, -1)
    c = np.corrcoef(flat)
    assert np.abs(c - np.eye(9)).max() < 2e-2                                     # across (step, utterance)
    for lag in (1, 2, 3, 4):                                                      # within a Philox block and across
        assert abs(np.corrcoef(flat[0, :-lag], flat[0, lag:])[0, 1]) < 2e-2
    # a pure function of (seed, step, utterance, sample): a batch offset or another batch size changes nothing
    assert np.array_equal(philox.noise(20240531, 3, 1, 1000, b_off=2)[:, 0], z[:, 2, :1000])
    assert not np.array_equal(philox.noise(20240532, 1, 1, 1000)[0, 0], z[0, 0, :1000])
