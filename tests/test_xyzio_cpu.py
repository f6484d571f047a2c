"""The on-disk format (host code of the C-ABI library, no GPU): pf_xyz_read / pf_xyz_write against numpy's
np.loadtxt(dtype=float32) / np.savetxt(fmt='%.6f'), which is what the reference uses (upsample.py:42,57)."""
import numpy as np
import pytest

from puflow_b200.xyzio import load_xyz, save_xyz
from puflow_b200._native import PuflowError


def _clouds():
    rng = np.random.default_rng(5)
    yield rng.standard_normal((2048, 3)).astype(np.float32)
    yield (rng.standard_normal((777, 3)) * 1e3).astype(np.float32)
    yield (rng.standard_normal((64, 6)) * 1e-4).astype(np.float32)           # xyz + normals
    yield (rng.standard_normal((4096, 3)) * np.exp(rng.uniform(-20, 20, (4096, 3)))).astype(np.float32)   # all magnitudes
    half = (np.arange(3000, dtype=np.float64).reshape(-1, 3) * 2 + 1) * 0.5e-6                             # exact ties x.xxxxxx5
    yield half.astype(np.float32)
    yield np.array([[0.0, -0.0, 1.0], [0.5e-6, -0.5e-6, 1.5e-6], [123456.789, -1e-7, 2.5e-6], [3.4e38, -3.4e38, 1e-38]], np.float32)


@pytest.mark.parametrize("k", range(6))
def test_write_is_byte_identical_to_savetxt(tmp_path, k):
    a = list(_clouds())[k]
    np.savetxt(tmp_path / "np.xyz", a, fmt="%.6f")
    save_xyz(tmp_path / "pf.xyz", a)
    assert (tmp_path / "np.xyz").read_bytes() == (tmp_path / "pf.xyz").read_bytes()


@pytest.mark.parametrize("k", range(6))
def test_read_is_bit_identical_to_loadtxt(tmp_path, k):
    a = list(_clouds())[k]
    np.savetxt(tmp_path / "six.xyz", a, fmt="%.6f")
    np.savetxt(tmp_path / "full.xyz", a, fmt="%.17g")
    np.savetxt(tmp_path / "exp.xyz", a, fmt="%.9e", delimiter="\t")
    for name in ("six.xyz", "full.xyz", "exp.xyz"):
        ref = np.loadtxt(tmp_path / name, dtype=np.float32)
        got = load_xyz(tmp_path / name)
        assert got.dtype == np.float32 and got.shape == ref.shape
        assert np.array_equal(got.view(np.uint32), ref.view(np.uint32)), name


def test_comments_blank_lines_and_errors(tmp_path):
    p = tmp_path / "c.xyz"
    p.write_text("# header\n1 2 3\n\n  4.5\t-6e-1   7 # trailing comment\r\n8 9 10")
    assert np.array_equal(load_xyz(p), np.loadtxt(p, dtype=np.float32))
    (tmp_path / "empty.xyz").write_text("")
    assert load_xyz(tmp_path / "empty.xyz").size == 0
    (tmp_path / "ragged.xyz").write_text("1 2 3\n4 5\n")
    with pytest.raises(PuflowError):
        load_xyz(tmp_path / "ragged.xyz")
    (tmp_path / "bad.xyz").write_text("1 2 x\n")
    with pytest.raises(PuflowError):
        load_xyz(tmp_path / "bad.xyz")
    with pytest.raises(PuflowError):
        load_xyz(tmp_path / "missing.xyz")


def test_round_trip_and_speed(tmp_path):
    import time
    a = np.random.default_rng(1).standard_normal((8192, 3)).astype(np.float32)
    t0 = time.perf_counter(); save_xyz(tmp_path / "a.xyz", a); b = load_xyz(tmp_path / "a.xyz"); t_pf = time.perf_counter() - t0
    t0 = time.perf_counter(); np.savetxt(tmp_path / "b.xyz", a, fmt="%.6f"); c = np.loadtxt(tmp_path / "b.xyz", dtype=np.float32); t_np = time.perf_counter() - t0
    assert np.array_equal(b, c) and np.abs(b - a).max() <= 5.1e-7 * max(1.0, np.abs(a).max())
    print(f"xyz round trip of 8,192 points: puflow_b200 {t_pf * 1e3:.1f} ms, numpy {t_np * 1e3:.1f} ms")
    assert t_pf < t_np
