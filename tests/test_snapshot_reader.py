"""the CONCEPT snapshot reader (nbody_b200/snapshot.py, no h5py) on a raw
snapshot file the reference ships (tests/golden/snapshot_test1.hdf5, copied by
oracle/gen_golden.py).  The same positions produce the reference's own shipped
delta_test_id-1.npy (test_oracle_golden.py), which pins the reader's byte order, shape and
dataset lookup; here the file interface itself is covered: attributes, momenta, the directory
walk of FraP_kTilDelta.py:30-41 and the error cases."""
import os
import shutil

import numpy as np
import pytest

from nbody_b200 import snapshot

GOLDEN = os.path.join(os.path.dirname(__file__), "golden")
SNAP = os.path.join(GOLDEN, "snapshot_test1.hdf5")


def test_positions_and_boxsize_match_the_golden_arrays():
    g = np.load(os.path.join(GOLDEN, "snap_test1.npz"))
    pos = snapshot.read_positions(SNAP)
    assert pos.dtype == np.float64 and pos.shape == (32 ** 3, 3) and pos.flags.c_contiguous
    assert np.array_equal(pos, g["pos"])
    box = snapshot.read_boxsize(SNAP)
    assert isinstance(box, float) and box == float(g["boxsize"]) == 500.0
    assert pos.min() >= 0.0 and pos.max() < box


def test_attributes_and_momenta():
    root = snapshot.read_attrs(SNAP)
    assert root["a"] == 1.0 and root["boxsize"] == 500.0
    assert {"H0", "Ωb", "Ωcdm"} <= set(root)
    comp = snapshot.read_attrs(SNAP, "components/matter")
    assert int(comp["N"]) == 32 ** 3 and comp["mass"] > 0
    mom = snapshot.read_momenta(SNAP)
    assert mom.shape == (32 ** 3, 3) and mom.dtype == np.float64 and np.isfinite(mom).all()
    # a relaxed snapshot: net momentum is small against the typical particle momentum
    assert np.abs(mom.mean(axis=0)).max() < 0.05 * np.abs(mom).mean()


def test_directory_walk_is_sorted_and_recursive(tmp_path):
    for rel in ("b/x/snapshots", "a/y/snapshots", "a/z"):
        os.makedirs(tmp_path / rel)
    shutil.copyfile(SNAP, tmp_path / "b/x/snapshots/snapshot_a=1.00.hdf5")
    shutil.copyfile(SNAP, tmp_path / "a/y/snapshots/snapshot_a=1.00.hdf5")
    (tmp_path / "a/z/notes.txt").write_text("not a snapshot")
    files = snapshot.find_hdf5_files(str(tmp_path))
    assert [os.path.relpath(f, tmp_path) for f in files] == [
        "a/y/snapshots/snapshot_a=1.00.hdf5", "b/x/snapshots/snapshot_a=1.00.hdf5"]
    assert snapshot.find_hdf5_files(str(tmp_path / "a/z")) == []


def test_error_cases(tmp_path):
    with pytest.raises(FileNotFoundError):
        snapshot.read_positions(str(tmp_path / "missing.hdf5"))
    bad = tmp_path / "bad.hdf5"
    bad.write_bytes(b"not an hdf5 file" * 64)
    with pytest.raises(ValueError):
        snapshot.read_positions(str(bad))
    with pytest.raises((KeyError, ValueError)):
        snapshot._H5(SNAP).dataset("components/matter/nope")
    # a truncated file must raise, not return garbage
    cut = tmp_path / "cut.hdf5"
    cut.write_bytes(open(SNAP, "rb").read()[:200000])
    with pytest.raises((ValueError, IndexError, EOFError, OSError)):
        snapshot.read_positions(str(cut))


def test_writer_round_trip_and_layout_matches_the_reference_file(tmp_path):
    """snapshot.write_snapshot: same object tree, attribute names, dataset
    shapes / types and dataset message encodings as the CONCEPT file; values bit-exact after a
    round trip.  (No libhdf5 in this image: the reader is the only consumer tested.)"""
    rs = np.random.RandomState(2)
    n = 4321                                       # not a multiple of anything convenient
    pos = rs.random_sample((n, 3)) * 500.0
    mom = rs.normal(scale=1e4, size=(n, 3))
    out = str(tmp_path / "snapshot_a=0.50.hdf5")
    ref_root = snapshot.read_attrs(SNAP)
    snapshot.write_snapshot(out, pos, mom, boxsize=500.0, a=0.5, mass=15072.97, H0=ref_root["H0"],
                            Omega_b=ref_root["Ωb"], Omega_cdm=ref_root["Ωcdm"])
    assert np.array_equal(snapshot.read_positions(out), pos)
    assert np.array_equal(snapshot.read_momenta(out), mom)
    assert snapshot.read_boxsize(out) == 500.0
    got_root = snapshot.read_attrs(out)
    assert set(ref_root) <= set(got_root) and got_root["a"] == 0.5
    for key in ("H0", "Ωb", "Ωcdm"):
        assert got_root[key] == ref_root[key]
    comp = snapshot.read_attrs(out, "components/matter")
    assert int(comp["N"]) == n and comp["mass"] == 15072.97 and comp["species"] == "matter"
    mine, ref = snapshot._H5(out), snapshot._H5(SNAP)
    assert set(mine.children(mine.root_header)) == set(ref.children(ref.root_header)) == {"components"}
    for h in (mine, ref):
        assert set(h.children(h.resolve("components/matter"))) == {"pos", "mom"}
    # dataset headers: identical datatype and fill-value message bytes, same message kinds
    for name in ("pos", "mom"):
        m = {t: b for t, _f, b in mine.messages(mine.resolve(f"components/matter/{name}"))}
        r = {t: b for t, _f, b in ref.messages(ref.resolve(f"components/matter/{name}"))}
        assert m[0x03] == r[0x03] and m[0x05] == r[0x05]
        assert m[0x01][:8] == r[0x01][:8] and m[0x08][:2] == r[0x08][:2]
    with pytest.raises(ValueError):
        snapshot.write_snapshot(out, pos, mom[:-1], boxsize=500.0)
    # the file on disk is complete: end-of-file address in the superblock equals its size
    import struct
    assert struct.unpack_from("<Q", open(out, "rb").read(96), 40)[0] == os.path.getsize(out)
