"""LoadOp / SaveOp edge: .npy and .csv files into device tensors (moose/src/storage/filesystem/numpy.rs, csv.rs),
with the fixed-point encoding fused into the load; checked against numpy's own reader / writer and the oracle's
encoder."""
import numpy as np
import pytest

import moose_b200 as mb
from moose_b200 import RING64, RING128
from oracle import moose_oracle as mo
from _util import ENTROPY_KEY, SESSION_ID, same_ring

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def sess():
    s = mb.Session(0, SESSION_ID, ENTROPY_KEY)
    yield s
    s.close()


@pytest.mark.parametrize("dtype", ["<f8", "<f4", "<i8", "<i4", "<u8", "<u4"])
def test_load_npy_types(sess, tmp_path, dtype):
    rng = np.random.default_rng(7)
    a = (rng.uniform(-1000, 1000, (37, 5, 3)) if "f" in dtype else rng.integers(0, 1 << 20, (37, 5, 3))).astype(dtype)
    path = str(tmp_path / "x.npy")
    np.save(path, a)
    got = sess.load_npy(path)
    assert got.shape == a.shape and np.array_equal(got.to_host(), a.astype(np.float64))


@pytest.mark.parametrize("bits,frac", [(64, 23), (128, 40)])
def test_load_npy_fixed_streams_several_chunks(sess, tmp_path, bits, frac):
    """9.4M float64 values (75 MB): three staging chunks; the fused encoding equals encode(load(file))."""
    a = np.random.default_rng(8).uniform(-4, 4, (3000, 3137))
    path = str(tmp_path / "big.npy")
    np.save(path, a)
    ring = RING64 if bits == 64 else RING128
    got = sess.load_npy(path, ring, frac)
    assert same_ring(got, mo.fixed_encode(bits, a, frac))
    assert np.array_equal(sess.load_npy(path).to_host(), a)


def test_save_npy_and_errors(sess, tmp_path):
    a = np.random.default_rng(9).normal(size=(11, 7))
    path = str(tmp_path / "y.npy")
    sess.save_npy(sess.from_host(mb.FLOAT64, a), path)
    assert np.array_equal(np.load(path), a)
    r = np.arange(12, dtype=np.uint64).reshape(3, 4)
    sess.save_npy(sess.from_host(RING64, r), path)
    back = np.load(path)
    assert back.dtype == np.uint64 and np.array_equal(back, r)
    sess.save_npy(sess.from_host(mb.FLOAT64, np.arange(5.0)), path)
    assert np.array_equal(np.load(path), np.arange(5.0))
    with pytest.raises(mb.MooseB200Error):
        sess.load_npy(str(tmp_path / "missing.npy"))
    np.save(path, np.asfortranarray(a))
    with pytest.raises(mb.MooseB200Error):
        sess.load_npy(path)
    np.save(path, a.astype(">f8"))
    with pytest.raises(mb.MooseB200Error):
        sess.load_npy(path)


def test_csv_roundtrip_and_column_selection(sess, tmp_path):
    path = str(tmp_path / "d.csv")
    with open(path, "w") as f:
        f.write("a,b,c\n1.5,2,3\n-4,5e-1,6\n")
    assert sess.load_csv(path).to_host().tolist() == [[1.5, 2.0, 3.0], [-4.0, 0.5, 6.0]]
    assert sess.load_csv(path, ["c", "a"]).to_host().tolist() == [[1.5, 3.0], [-4.0, 6.0]]  # file order, like read_csv
    out = str(tmp_path / "o.csv")
    a = np.random.default_rng(10).normal(size=(6, 4))
    sess.save_csv(sess.from_host(mb.FLOAT64, a), out)
    assert open(out).readline().strip() == "col_0,col_1,col_2,col_3"
    assert np.array_equal(sess.load_csv(out).to_host(), a)
    with open(path, "w") as f:
        f.write("a,b\n1,x\n")
    with pytest.raises(mb.MooseB200Error):
        sess.load_csv(path)
