"""CPU-side tests: C-ABI exports, COLMAP database layout, pair generation / sharding."""
import ctypes
import re
import sqlite3
import struct
from pathlib import Path

import numpy as np
import pytest

ROOT = Path(__file__).resolve().parent.parent


def test_c_abi_exports_every_declared_symbol():
    from vit_colmap_b200 import _lib, build
    build.build()
    header = (ROOT / "include" / "vc_b200.h").read_text()
    declared = set(re.findall(r"\b(vc_[a-z0-9_]+)\s*\(", header))
    assert declared, "no declarations parsed"
    lib = ctypes.CDLL(str(_lib.LIB_PATH))
    for name in declared:
        assert hasattr(lib, name), f"{name} declared in vc_b200.h but not exported"
    assert declared == set(_lib.SIGNATURES), "ctypes table out of sync with the header"
    assert b"sm_100a" in _lib.load().vc_version()


def test_c_abi_header_is_plain_c(tmp_path):
    """include/vc_b200.h is what a C / cgo / JNI binding would include: it must compile as C99 on its own."""
    import shutil
    import subprocess
    if shutil.which("gcc") is None:
        pytest.skip("no gcc")
    src = tmp_path / "t.c"
    src.write_text('#include "vc_b200.h"\nint main(void) { return 0; }\n')
    r = subprocess.run(["gcc", "-std=c99", "-Wall", "-Wextra", "-pedantic", "-Werror", "-fsyntax-only",
                        "-I", str(ROOT / "include"), str(src)], capture_output=True, text=True)
    assert r.returncode == 0, r.stderr


def test_no_device_fails_loudly():
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    from vit_colmap_b200 import _lib
    from vit_colmap_b200.matching import Matcher
    with pytest.raises(_lib.VcError):
        Matcher(0)
    h = ctypes.c_void_p()
    rc = _lib.load().vc_create(0, ctypes.byref(h))
    assert rc == -3 and b"no CPU fallback" in _lib.load().vc_last_error()


def test_database_layout(tmp_path):
    from vit_colmap_b200.database import ColmapDatabase, make_camera
    from vit_colmap_b200.pairing import image_pair_to_pair_id
    path = tmp_path / "db.db"
    cdb = ColmapDatabase(str(path))
    cam = cdb.db.write_camera(make_camera("PINHOLE", 640, 480, [640.0, 640.0, 320.0, 240.0]))
    ids = [cdb.add_image(f"im{i}.png", cam) for i in range(3)]
    assert ids == [1, 2, 3]
    kp = np.arange(12, dtype=np.float64).reshape(6, 2)
    desc = (np.arange(6 * 128) % 251).reshape(6, 128)
    cdb.add_keypoints(1, kp)
    cdb.add_descriptors(1, desc)
    m = np.array([[0, 5], [3, 1]], dtype=np.int64)
    cdb.add_matches(2, 1, m)          # id1 > id2 -> stored swapped (database_sqlite.cc:1115-1121)
    cdb.add_matches(1, 3, np.zeros((0, 2)))
    assert cdb.get_db_count(cdb.db, "num_images") == 3
    assert cdb.get_db_count(cdb.db, "num_cameras") == 1
    assert cdb.get_db_count(cdb.db, "num_matched_image_pairs") == 2  # empty rows count
    assert cdb.db.exists_keypoints(1) and cdb.db.exists_descriptors(1) and not cdb.db.exists_keypoints(2)
    assert cdb.db.read_matches(2, 1).tolist() == m.tolist()
    assert cdb.db.read_matches(1, 2).tolist() == m[:, ::-1].tolist()
    cdb.db.close()
    con = sqlite3.connect(str(path))
    assert con.execute("PRAGMA user_version").fetchone()[0] == 3900
    tables = {r[0] for r in con.execute("SELECT name FROM sqlite_master WHERE type='table'")}
    assert {"cameras", "images", "keypoints", "descriptors", "matches", "two_view_geometries", "rigs",
            "rig_sensors", "frames", "frame_data", "pose_priors"} <= tables
    rows, cols, data = con.execute("SELECT rows, cols, data FROM keypoints WHERE image_id=1").fetchone()
    assert (rows, cols) == (6, 2) and data == kp.astype(np.float32).tobytes()
    rows, cols, data = con.execute("SELECT rows, cols, data FROM descriptors WHERE image_id=1").fetchone()
    assert (rows, cols) == (6, 128) and data == desc.astype(np.uint8).tobytes()
    pid = image_pair_to_pair_id(2, 1)
    assert pid == 2147483647 * 1 + 2
    rows, cols, data = con.execute("SELECT rows, cols, data FROM matches WHERE pair_id=?", (pid,)).fetchone()
    assert (rows, cols) == (2, 2) and struct.unpack("<4I", data) == (5, 0, 1, 3)
    model, w, h, params, prior = con.execute("SELECT model,width,height,params,prior_focal_length FROM cameras").fetchone()
    assert (model, w, h, prior) == (1, 640, 480, 0) and struct.unpack("<4d", params) == (640.0, 640.0, 320.0, 240.0)
    with ColmapDatabase.open_database(str(path)) as db:
        assert ColmapDatabase.get_db_count(db, "num_images") == 3


def test_pair_generator_matches_oracle_and_shards():
    from oracle import matching as om
    from vit_colmap_b200.pairing import ExhaustivePairGenerator, shard_blocks, shard_pairs
    for n, bs in [(34, 10), (49, 50), (8, 3), (1, 50), (0, 50), (101, 50)]:
        ids = list(range(1, n + 1))
        gen = ExhaustivePairGenerator(ids, bs)
        blocks = list(gen.blocks())
        ref, sizes = om.exhaustive_pairs(np.array(ids, np.uint32), bs) if n else (np.zeros((0, 2)), [])
        flat = [p for b in blocks for p in b]
        assert flat == [tuple(x) for x in np.asarray(ref).tolist()]
        assert [len(b) for b in blocks] == list(sizes)
        assert len({(min(a, b), max(a, b)) for a, b in flat}) == n * (n - 1) // 2
        for world in (1, 2, 8):
            assign = shard_blocks([len(b) for b in blocks], world)
            assert sorted(x for a in assign for x in a) == list(range(len(blocks)))
            parts = [shard_pairs(flat, r, world) for r in range(world)]
            assert [p for part in parts for p in part] == flat


def test_two_view_geometry_rows_follow_colmap_layout(tmp_path):
    """two_view_geometries rows (database_sqlite.cc:1128-1190, :839-881): stored for the (smaller id,
    larger id) order with the geometry inverted when written the other way round (F, E transposed,
    H inverted, pose inverted, match columns swapped), matrices as row-major float64 blobs, no matrix
    blobs without inlier matches."""
    import sqlite3
    import numpy as np
    from vit_colmap_b200.database import Database, TwoViewGeometry, image_pair_to_pair_id
    path = tmp_path / "t.db"
    F = np.arange(9.0).reshape(3, 3)
    H = np.array([[2.0, 0.0, 1.0], [0.0, 4.0, 0.0], [0.0, 0.0, 1.0]])
    q = np.array([np.cos(0.3), np.sin(0.3), 0.0, 0.0])
    g = TwoViewGeometry(3, F, F + 1, H, q, np.array([1.0, 2.0, 3.0]), np.array([[1, 2], [3, 4]], np.uint32))
    with Database.open(str(path)) as db:
        db.write_two_view_geometry(5, 2, g)
        db.write_two_view_geometry(2, 7, TwoViewGeometry(1))
        back = db.read_two_view_geometry(5, 2)
        assert np.allclose(back.F, F) and np.allclose(back.H, H) and back.inlier_matches.tolist() == [[1, 2], [3, 4]]
        assert np.allclose(back.qvec, q) and np.allclose(back.tvec, [1, 2, 3])
        fwd = db.read_two_view_geometry(2, 5)
        assert np.allclose(fwd.F, F.T) and np.allclose(fwd.H, np.linalg.inv(H)) and fwd.inlier_matches.tolist() == [[2, 1], [4, 3]]
        assert db.num_verified_image_pairs() == 2 and db.exists_two_view_geometry(7, 2)
        assert [pr for pr, _ in db.read_two_view_geometries()] == [(2, 5), (2, 7)]
    con = sqlite3.connect(str(path))
    rows, cols, data, config, fb, eb, hb, qb, tb = con.execute(
        "SELECT rows, cols, data, config, F, E, H, qvec, tvec FROM two_view_geometries WHERE pair_id = ?",
        (image_pair_to_pair_id(2, 5),)).fetchone()
    assert (rows, cols, config) == (2, 2, 3)
    assert np.frombuffer(data, np.uint32).tolist() == [2, 1, 4, 3]
    assert np.frombuffer(fb, np.float64).tolist() == F.T.reshape(-1).tolist()      # row-major
    assert len(hb) == 72 and len(qb) == 32 and len(tb) == 24 and len(eb) == 72
    empty = con.execute("SELECT rows, F, H, qvec FROM two_view_geometries WHERE pair_id = ?",
                        (image_pair_to_pair_id(2, 7),)).fetchone()
    assert empty[0] == 0 and all(not b for b in empty[1:])


def test_decode_ahead_keeps_order_and_passes_failures_through():
    """features/_pipelined.decode_ahead: results come back in input order (the single database writer
    hands out image ids in sorted file order), a failing prepare() is returned in place of its result,
    None (unreadable file) passes through, and no more than `lookahead` files are prepared ahead."""
    import threading
    import time
    from vit_colmap_b200.features._pipelined import decode_ahead
    started, lock = [], threading.Lock()

    def prepare(i):
        with lock:
            started.append(i)
        time.sleep(0.002 * ((i * 7) % 5))      # out-of-order completion
        if i == 3:
            raise ValueError("too small")
        return None if i == 5 else i * i

    ahead = []
    out = []
    for k, item, got in decode_ahead(range(30), prepare, workers=4, lookahead=6):
        with lock:
            ahead.append(max(started) - k)
        out.append((k, item, got))
    assert [k for k, _, _ in out] == list(range(30)) and [i for _, i, _ in out] == list(range(30))
    assert isinstance(out[3][2], ValueError) and out[5][2] is None
    assert [g for k, _, g in out if k not in (3, 5)] == [i * i for i in range(30) if i not in (3, 5)]
    assert max(ahead) <= 6
    assert list(decode_ahead([], prepare)) == []


def test_pad_descriptors_layout():
    """matching.pad_descriptors: ragged lists -> one padded buffer; rows beyond each count and columns beyond
    the descriptor width are zero (a width that is not a multiple of 8 is padded up), the row clamp follows
    max_num_matches (SIFTGPU/SiftMatchCU.cpp:108), a staging list is reused when large enough."""
    import numpy as np
    from vit_colmap_b200.matching import pad_descriptors
    rng = np.random.default_rng(0)
    descs = [rng.integers(1, 256, (5, 100), dtype=np.uint8), np.zeros((0, 100), np.uint8),
             rng.integers(1, 256, (9, 100), dtype=np.uint8)]
    staging = []
    buf, counts = pad_descriptors(descs, 7, pinned=False, staging=staging)
    a = buf.numpy()
    assert a.shape == (3, 7, 104) and counts.tolist() == [5, 0, 7]
    assert np.array_equal(a[0, :5, :100], descs[0]) and not a[0, 5:].any() and not a[:, :, 100:].any()
    assert not a[1].any() and np.array_equal(a[2, :, :100], descs[2][:7])
    first = staging[0]
    staging[0][:] = 255                                       # stale bytes from an earlier, larger upload
    buf2, counts2 = pad_descriptors([d.astype(np.float32) for d in descs[:1]], 32768, pinned=False, staging=staging)
    assert staging[0] is first and buf2.dtype.is_floating_point and tuple(buf2.shape) == (1, 5, 104)
    assert np.array_equal(buf2.numpy()[0, :, :100], descs[0].astype(np.float32)) and not buf2.numpy()[0, :, 100:].any()
