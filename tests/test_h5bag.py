"""CPU suite: the feature-bag file reader (reference reads at code/datasets/feature_dataloader.py:252-254).
Pinned two ways: (1) a file written by the real HDF5 library (scipy ships it as MATLAB v7.3 test data: 512-byte user block,
superblock 0, symbol-table group, version-1 object header, contiguous float64) whose content scipy documents as
0:pi/4:2*pi; (2) files written by tests/h5_writer.py in the layout the reference's converters produce (chunks=(1, D),
resizable; code/datasets/utils/move_to_hdf5.py:19) and its variations."""
import os

import numpy as np
import pytest
import torch

import h5_writer as W
from transmil_b200 import h5bag


def _real_hdf5_file():
    try:
        import scipy.io
    except ImportError:
        return None
    p = os.path.join(os.path.dirname(scipy.io.__file__), "matlab", "tests", "data", "testhdf5_7.4_GLNX86.mat")
    return p if os.path.isfile(p) else None


@pytest.mark.skipif(_real_hdf5_file() is None, reason="scipy's HDF5 test file is not installed")
def test_reads_a_file_written_by_the_hdf5_library():
    with h5bag.FeatureFile(_real_hdf5_file()) as f:
        assert f.keys() == ["testdouble"] and "testdouble" in f
        ds = f["testdouble"]
        assert ds.shape == (9, 1) and ds.dtype == np.dtype("<f8") and ds.chunks is None
        assert np.array_equal(ds[:].ravel(), np.arange(0, 2 * np.pi + np.pi / 4, np.pi / 4)[:9])


def _bag(n=777, D=96, dtype=np.float32, seed=0):
    rng = np.random.default_rng(seed)
    return rng.random((n, D)).astype(dtype), rng.integers(0, 200000, (n, 2)).astype(np.int64)


VARIANTS = {
    "contiguous": dict(),
    "reference-converter": dict(chunks=(1, 96), maxshape=(None, 96)),
    "deep-btree-interleaved": dict(chunks=(1, 96), maxshape=(None, 96), btree_cap=4, interleave=True, split_header=True),
    "rows-of-7-shuffle-deflate": dict(chunks=(7, 96), shuffle=True, deflate=4, btree_cap=5),
    "tiles-33x10": dict(chunks=(33, 10), btree_cap=3),
    "deflate-only": dict(chunks=(64, 96), deflate=1),
}


@pytest.mark.parametrize("variant", list(VARIANTS))
@pytest.mark.parametrize("userblock", [0, 512])
def test_round_trip_layouts(tmp_path, variant, userblock):
    x, c = _bag()
    p = W.write_h5(tmp_path / "slide.h5", {"features": (x, VARIANTS[variant]), "coords": c}, userblock=userblock)
    with h5bag.FeatureFile(p) as f:
        assert sorted(f.keys()) == ["coords", "features"]
        ds = f["features"]
        assert ds.shape == x.shape and ds.dtype == x.dtype and len(ds) == len(x)
        assert ds.chunks == VARIANTS[variant].get("chunks")
        assert np.array_equal(ds[:], x) and np.array_equal(ds[...], x) and np.array_equal(ds[10:20], x[10:20])
        assert np.array_equal(f["coords"][:], c)
        with pytest.raises(KeyError):
            f["nope"]


@pytest.mark.parametrize("dtype", [np.float16, np.float32, np.float64, np.int32, np.uint8, ">f4"])
def test_element_types(tmp_path, dtype):
    x = (np.random.default_rng(1).random((50, 8)) * 100).astype(dtype)
    p = W.write_h5(tmp_path / "t.h5", {"features": (x, dict(chunks=(1, 8)))})
    with h5bag.FeatureFile(p) as f:
        got = f["features"][:]
    assert got.dtype.itemsize == x.dtype.itemsize and np.array_equal(got, x)


def test_unallocated_chunks_read_as_fill_value_and_empty_dataset(tmp_path):
    x, _ = _bag(40, 8)
    p = W.write_h5(tmp_path / "s.h5", {"features": (x, dict(chunks=(4, 8), skip_chunks=((2, 0), (9, 0)))),
                                       "empty": (np.zeros((0, 8), np.float32), dict(chunks=(1, 8), maxshape=(None, 8)))})
    with h5bag.FeatureFile(p) as f:
        got = f["features"][:]
        assert f["empty"][:].shape == (0, 8)
    want = x.copy()
    want[8:12] = 0
    want[36:40] = 0
    assert np.array_equal(got, want)
    p = W.write_h5(tmp_path / "f.h5", {"features": (x, dict(chunks=(4, 8), skip_chunks=((2, 0),), fillvalue=-1.5))})
    with h5bag.FeatureFile(p) as f:
        assert f["features"].fillvalue == np.float32(-1.5)
        got = f["features"][:]
    want = x.copy()
    want[8:12] = -1.5
    assert np.array_equal(got, want)


def test_read_bag_fills_a_caller_owned_host_tensor(tmp_path):
    x, c = _bag(300, 64, np.float32, 3)
    p = W.write_h5(tmp_path / "b.h5", {"features": (x, dict(chunks=(1, 64), maxshape=(None, 64))), "coords": c})
    bag, xy = h5bag.read_bag(p, pin=False)                       # feature_dataloader.py:252-254,273
    assert bag.dtype == torch.float32 and torch.equal(bag, torch.from_numpy(x)) and torch.equal(xy, torch.from_numpy(c))
    half, _ = h5bag.read_bag(p, pin=False, dtype=torch.float16)  # converted while reading
    assert half.dtype == torch.float16 and torch.equal(half, torch.from_numpy(x).half())
    x16 = x.astype(np.float16)
    p16 = W.write_h5(tmp_path / "h.h5", {"features": (x16, dict(chunks=(1, 64)))})
    bag16, none = h5bag.read_bag(p16, pin=False)                 # fp16 feature files stay fp16 (fc1 reads halves), no coords
    assert bag16.dtype == torch.float16 and none is None and torch.equal(bag16, torch.from_numpy(x16))
    staging = torch.empty(1 << 16, dtype=torch.float32)          # a reused (pinned, on the GPU box) staging buffer
    view, _ = h5bag.read_bag(p, out=staging)
    assert view.data_ptr() == staging.data_ptr() and torch.equal(view, torch.from_numpy(x))
    with pytest.raises(ValueError):
        h5bag.read_bag(p, out=torch.empty(10))
    bags = list(h5bag.iter_bags([p, p16], pin=False))
    assert [b.dtype for b in bags] == [torch.float32, torch.float16]
    with h5bag.FeatureFile(p) as f:
        with pytest.raises(ValueError):
            f["features"].read_direct(np.empty((3, 3), np.float32))


def test_rejects_what_it_does_not_implement(tmp_path):
    bad = tmp_path / "x.h5"
    bad.write_bytes(b"not an hdf5 file at all" * 10)
    with pytest.raises(h5bag.H5FormatError, match="signature"):
        h5bag.FeatureFile(bad)
    x, _ = _bag(10, 4)
    p = W.write_h5(tmp_path / "v.h5", {"features": x})
    raw = bytearray(open(p, "rb").read())
    raw[8] = 2                                                   # superblock version 2 (libver='latest')
    (tmp_path / "v2.h5").write_bytes(raw)
    with pytest.raises(h5bag.H5FormatError, match="superblock version 2"):
        h5bag.FeatureFile(tmp_path / "v2.h5")
    with pytest.raises(ValueError):
        h5bag.FeatureFile(p, "w")
    trunc = tmp_path / "t.h5"
    trunc.write_bytes(bytes(raw[:200]).replace(b"\x02", b"\x00", 1))
    with pytest.raises(h5bag.H5FormatError):
        with h5bag.FeatureFile(trunc) as f:
            f["features"][:]


def test_reference_sized_bag_reads_in_few_copies(tmp_path):
    """20 000 x 2048 would be 164 MB; 20 000 x 64 keeps the chunk count (the B-tree depth and the run coalescing) and the test short."""
    x, _ = _bag(20000, 64, np.float32, 5)
    p = W.write_h5(tmp_path / "big.h5", {"features": (x, dict(chunks=(1, 64), maxshape=(None, 64), btree_cap=64))})
    bag, _ = h5bag.read_bag(p, pin=False)
    assert torch.equal(bag, torch.from_numpy(x))


def test_threaded_prefetch_keeps_order_and_content(tmp_path):
    paths, arrays = [], []
    for i, n in enumerate([40, 1, 300, 7, 120, 64, 5, 200, 33]):
        x, c = _bag(n, 32, np.float32, 20 + i)
        paths.append(W.write_h5(tmp_path / f"s{i}.h5", {"features": (x, dict(chunks=(1, 32), maxshape=(None, 32))), "coords": c}))
        arrays.append((x, c))
    for workers in (0, 1, 3):
        got = list(h5bag.iter_bags(paths, pin=False, workers=workers, with_coords=True))
        assert len(got) == len(arrays)
        for (bag, xy), (x, c) in zip(got, arrays):
            assert torch.equal(bag, torch.from_numpy(x)) and torch.equal(xy, torch.from_numpy(c))
    it = h5bag.iter_bags(paths + [tmp_path / "missing.h5"], pin=False, workers=2)
    with pytest.raises(FileNotFoundError):
        list(it)


@pytest.mark.parametrize("variant", list(VARIANTS))
def test_row_ranges_touch_only_their_chunks(tmp_path, variant):
    x, _ = _bag(501, 96, np.float32, 9)
    p = W.write_h5(tmp_path / "r.h5", {"features": (x, VARIANTS[variant])})
    with h5bag.FeatureFile(p) as f:
        ds = f["features"]
        for r0, r1 in [(0, 501), (0, 1), (500, 501), (13, 77), (66, 99), (7, 7), (33, 330), (495, 501)]:
            assert np.array_equal(ds.read_rows(r0, r1), x[r0:r1]), (r0, r1)
            assert np.array_equal(ds[r0:r1], x[r0:r1])
        assert np.array_equal(ds[-5:], x[-5:]) and np.array_equal(ds[::2], x[::2]) and np.array_equal(ds[3], x[3])
        with pytest.raises(IndexError):
            ds.read_rows(10, 502)
    ranges = [(400, 501), (0, 37)]
    got = h5bag.read_bag_rows(p, ranges, pin=False)
    assert torch.equal(got, torch.from_numpy(np.concatenate([x[a:b] for a, b in ranges])))


@pytest.mark.parametrize("world", [2, 4])
def test_shard_plan_reads_only_its_rows_of_a_file(tmp_path, world):
    """ShardPlan.take_from_file == ShardPlan.take on the full bag, for every rank of a token-sharded bag."""
    from transmil_b200 import sharded
    n, D = 5000, 32
    x, _ = _bag(n, D, np.float32, 10)
    p = W.write_h5(tmp_path / "giant.h5", {"features": (x, dict(chunks=(1, D), maxshape=(None, D)))})
    plan = sharded.ShardPlan(n, world)
    full = torch.from_numpy(x)
    for rank in range(world):
        assert torch.equal(plan.take_from_file(p, rank, pin=False), plan.take(full, rank))
