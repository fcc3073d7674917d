"""COLMAP database writer/reader on plain sqlite3, byte-compatible with COLMAP's schema.

Mirrors the reference wrapper ``vit_colmap/database/colmap_db.py:6-75`` (``ColmapDatabase``) and
the part of ``pycolmap.Database`` it and the tests use (PYCOLMAP/scene/database.cc:388-547).
On-disk contract restated from COLMAP/scene/database_sqlite.cc: pragmas :437-449, tables
:1811-1958, blob layouts :47-99,:192-207 (row-major rows/cols/data), pair ids and the column swap
for id1 > id2 :1099-1123, user_version 3900 :2013-2015.  When pycolmap is importable the reference
wrapper's behaviour (delegating to pycolmap.Database) is kept unchanged.
"""
from __future__ import annotations

import sqlite3
from contextlib import contextmanager
from dataclasses import dataclass, field
from typing import List, Optional, Sequence, Tuple

import numpy as np

from .pairing import image_pair_to_pair_id, pair_id_to_image_pair, K_MAX_NUM_IMAGES

try:  # pragma: no cover - not installed in the build image
    import pycolmap as _pycolmap
except Exception:  # noqa: BLE001
    _pycolmap = None

# COLMAP/sensor/models.h:84-99
CAMERA_MODEL_IDS = {
    "SIMPLE_PINHOLE": 0, "PINHOLE": 1, "SIMPLE_RADIAL": 2, "RADIAL": 3, "OPENCV": 4,
    "OPENCV_FISHEYE": 5, "FULL_OPENCV": 6, "FOV": 7, "SIMPLE_RADIAL_FISHEYE": 8,
    "RADIAL_FISHEYE": 9, "THIN_PRISM_FISHEYE": 10,
}
CAMERA_MODEL_NUM_PARAMS = {0: 3, 1: 4, 2: 4, 3: 5, 4: 8, 5: 8, 6: 12, 7: 5, 8: 4, 9: 5, 10: 12}

_SCHEMA = f"""
CREATE TABLE IF NOT EXISTS rigs
   (rig_id INTEGER PRIMARY KEY AUTOINCREMENT NOT NULL,
    ref_sensor_id INTEGER NOT NULL,
    ref_sensor_type INTEGER NOT NULL);
CREATE UNIQUE INDEX IF NOT EXISTS rig_ref_sensor_assignment ON rigs(ref_sensor_id, ref_sensor_type);
CREATE TABLE IF NOT EXISTS rig_sensors
   (rig_id INTEGER NOT NULL,
    sensor_id INTEGER NOT NULL,
    sensor_type INTEGER NOT NULL,
    sensor_from_rig BLOB,
    FOREIGN KEY(rig_id) REFERENCES rigs(rig_id) ON DELETE CASCADE);
CREATE UNIQUE INDEX IF NOT EXISTS rig_sensor_assignment ON rig_sensors(sensor_id, sensor_type);
CREATE TABLE IF NOT EXISTS cameras
   (camera_id INTEGER PRIMARY KEY AUTOINCREMENT NOT NULL,
    model INTEGER NOT NULL,
    width INTEGER NOT NULL,
    height INTEGER NOT NULL,
    params BLOB,
    prior_focal_length INTEGER NOT NULL);
CREATE TABLE IF NOT EXISTS frames
   (frame_id INTEGER PRIMARY KEY AUTOINCREMENT NOT NULL,
    rig_id INTEGER NOT NULL,
    FOREIGN KEY(rig_id) REFERENCES rigs(rig_id) ON DELETE CASCADE);
CREATE TABLE IF NOT EXISTS frame_data
   (frame_id INTEGER NOT NULL,
    data_id INTEGER NOT NULL,
    sensor_id INTEGER NOT NULL,
    sensor_type INTEGER NOT NULL,
    FOREIGN KEY(frame_id) REFERENCES frames(frame_id) ON DELETE CASCADE);
CREATE UNIQUE INDEX IF NOT EXISTS frame_sensor_assignment ON frame_data(data_id, sensor_type);
CREATE TABLE IF NOT EXISTS images
   (image_id INTEGER PRIMARY KEY AUTOINCREMENT NOT NULL,
    name TEXT NOT NULL UNIQUE,
    camera_id INTEGER NOT NULL,
    CONSTRAINT image_id_check CHECK(image_id >= 0 and image_id < {K_MAX_NUM_IMAGES}),
    FOREIGN KEY(camera_id) REFERENCES cameras(camera_id));
CREATE UNIQUE INDEX IF NOT EXISTS index_name ON images(name);
CREATE TABLE IF NOT EXISTS pose_priors
   (image_id INTEGER PRIMARY KEY NOT NULL,
    position BLOB,
    coordinate_system INTEGER NOT NULL,
    position_covariance BLOB,
    FOREIGN KEY(image_id) REFERENCES images(image_id) ON DELETE CASCADE);
CREATE TABLE IF NOT EXISTS keypoints
   (image_id INTEGER PRIMARY KEY NOT NULL,
    rows INTEGER NOT NULL,
    cols INTEGER NOT NULL,
    data BLOB,
    FOREIGN KEY(image_id) REFERENCES images(image_id) ON DELETE CASCADE);
CREATE TABLE IF NOT EXISTS descriptors
   (image_id INTEGER PRIMARY KEY NOT NULL,
    rows INTEGER NOT NULL,
    cols INTEGER NOT NULL,
    data BLOB,
    FOREIGN KEY(image_id) REFERENCES images(image_id) ON DELETE CASCADE);
CREATE TABLE IF NOT EXISTS matches
   (pair_id INTEGER PRIMARY KEY NOT NULL,
    rows INTEGER NOT NULL,
    cols INTEGER NOT NULL,
    data BLOB);
CREATE TABLE IF NOT EXISTS two_view_geometries
   (pair_id INTEGER PRIMARY KEY NOT NULL,
    rows INTEGER NOT NULL,
    cols INTEGER NOT NULL,
    data BLOB,
    config INTEGER NOT NULL,
    F BLOB,
    E BLOB,
    H BLOB,
    qvec BLOB,
    tvec BLOB);
"""


@dataclass
class Camera:
    """The fields of pycolmap.Camera the extractors set (vit_extractor.py:722-724)."""
    model: str = "SIMPLE_PINHOLE"
    width: int = 0
    height: int = 0
    params: Sequence[float] = field(default_factory=list)
    camera_id: int = -1
    has_prior_focal_length: bool = False

    @property
    def model_id(self) -> int:
        if self.model not in CAMERA_MODEL_IDS:
            raise ValueError(f"Unsupported camera model: {self.model}")
        return CAMERA_MODEL_IDS[self.model]


@dataclass
class Image:
    """pycolmap.Image(name=..., camera_id=...) as used by colmap_db.py:28-30."""
    name: str = ""
    camera_id: int = -1
    image_id: int = -1


@dataclass
class TwoViewGeometry:
    """COLMAP/scene/two_view_geometry.h:38-85 (the fields the database stores)."""
    config: int = 0  # ConfigurationType: 0 UNDEFINED ... 2 CALIBRATED, 3 UNCALIBRATED, 4-6 homographies
    F: np.ndarray = field(default_factory=lambda: np.zeros((3, 3)))
    E: np.ndarray = field(default_factory=lambda: np.zeros((3, 3)))
    H: np.ndarray = field(default_factory=lambda: np.zeros((3, 3)))
    qvec: np.ndarray = field(default_factory=lambda: np.array([1.0, 0.0, 0.0, 0.0]))  # cam2_from_cam1, wxyz
    tvec: np.ndarray = field(default_factory=lambda: np.zeros(3))
    inlier_matches: np.ndarray = field(default_factory=lambda: np.zeros((0, 2), np.uint32))

    def inverted(self) -> "TwoViewGeometry":
        """TwoViewGeometry::Invert (scene/two_view_geometry.cc:34-42): the same geometry seen from image 2."""
        w, x, y, z = (float(v) for v in self.qvec)
        rot = np.array([[1 - 2 * (y * y + z * z), 2 * (x * y - z * w), 2 * (x * z + y * w)],
                        [2 * (x * y + z * w), 1 - 2 * (x * x + z * z), 2 * (y * z - x * w)],
                        [2 * (x * z - y * w), 2 * (y * z + x * w), 1 - 2 * (x * x + y * y)]])
        H = np.asarray(self.H, dtype=np.float64)
        H_inv = np.linalg.inv(H) if abs(np.linalg.det(H)) > 0 else np.zeros((3, 3))
        return TwoViewGeometry(self.config, np.asarray(self.F).T.copy(), np.asarray(self.E).T.copy(), H_inv,
                               np.array([w, -x, -y, -z]), -rot.T @ np.asarray(self.tvec, dtype=np.float64),
                               np.asarray(self.inlier_matches, dtype=np.uint32).reshape(-1, 2)[:, ::-1].copy())


class Database:
    """sqlite3-backed stand-in for pycolmap.Database (methods used on the hot path)."""

    def __init__(self, path: Optional[str] = None):
        self._conn: Optional[sqlite3.Connection] = None
        self._in_tx = False
        if path is not None:
            self._open(path)

    # pycolmap 3.13: ``Database.open(path)`` static; 3.12: ``Database().open(path)``
    @staticmethod
    def open(path: str) -> "Database":
        return Database(str(path))

    def _open(self, path: str) -> None:
        conn = sqlite3.connect(str(path), isolation_level=None)
        # database_sqlite.cc:437-449
        # auto_vacuum must precede table creation to take effect on a fresh file
        conn.execute("PRAGMA auto_vacuum=1")
        conn.execute("PRAGMA synchronous=OFF")
        conn.execute("PRAGMA journal_mode=WAL")
        conn.execute("PRAGMA temp_store=MEMORY")
        conn.execute("PRAGMA foreign_keys=ON")
        conn.executescript(_SCHEMA)
        conn.execute("PRAGMA user_version = 3900")  # database_sqlite.cc:2013-2015
        self._conn = conn

    def close(self) -> None:
        if self._conn is not None:
            if self._in_tx:
                self._conn.execute("COMMIT")
                self._in_tx = False
            self._conn.close()
            self._conn = None

    def __enter__(self) -> "Database":
        return self

    def __exit__(self, *exc) -> None:
        self.close()

    @contextmanager
    def transaction(self):
        """Batch many writes into one transaction (COLMAP's DatabaseTransaction,
        used around the write loop in controllers/feature_matching_utils.cc:478)."""
        if self._in_tx:
            yield self
            return
        self._conn.execute("BEGIN")
        self._in_tx = True
        try:
            yield self
        except Exception:
            self._conn.execute("ROLLBACK")
            self._in_tx = False
            raise
        self._conn.execute("COMMIT")
        self._in_tx = False

    # ------------------------------------------------------------------ writes
    def write_camera(self, camera: Camera, use_camera_id: bool = False) -> int:
        params = np.asarray(list(camera.params), dtype=np.float64)
        expected = CAMERA_MODEL_NUM_PARAMS[camera.model_id]
        if params.size != expected:
            raise ValueError(f"{camera.model} expects {expected} params, got {params.size}")
        cur = self._conn.execute(
            "INSERT INTO cameras(camera_id, model, width, height, params, prior_focal_length) "
            "VALUES(?, ?, ?, ?, ?, ?)",
            (camera.camera_id if use_camera_id else None, camera.model_id, int(camera.width),
             int(camera.height), params.tobytes(), int(bool(camera.has_prior_focal_length))))
        return int(cur.lastrowid)

    def write_image(self, image: Image, use_image_id: bool = False) -> int:
        cur = self._conn.execute(
            "INSERT INTO images(image_id, name, camera_id) VALUES(?, ?, ?)",
            (image.image_id if use_image_id else None, image.name, int(image.camera_id)))
        return int(cur.lastrowid)

    def _write_blob(self, table: str, key: int, arr: np.ndarray) -> None:
        arr = np.ascontiguousarray(arr)
        rows, cols = (arr.shape if arr.ndim == 2 else (0, 0))
        keycol = "pair_id" if table == "matches" else "image_id"
        self._conn.execute(f"INSERT INTO {table}({keycol}, rows, cols, data) VALUES(?, ?, ?, ?)",
                           (int(key), int(rows), int(cols), arr.tobytes()))

    def write_keypoints(self, image_id: int, keypoints: np.ndarray) -> None:
        kp = np.asarray(keypoints, dtype=np.float32)
        if kp.ndim != 2 or kp.shape[1] not in (2, 4, 6):  # database_sqlite.cc:66-88
            raise ValueError("keypoints must be [N, 2|4|6]")
        self._write_blob("keypoints", image_id, kp)

    def write_descriptors(self, image_id: int, descriptors: np.ndarray) -> None:
        self._write_blob("descriptors", image_id, np.asarray(descriptors, dtype=np.uint8))

    def write_matches(self, image_id1: int, image_id2: int, matches: np.ndarray) -> None:
        m = np.asarray(matches, dtype=np.uint32).reshape(-1, 2)
        if image_id1 > image_id2:  # SwapImagePair: database_sqlite.cc:1115-1121
            m = m[:, ::-1]
        if m.shape[0] == 0:
            m = np.zeros((0, 2), np.uint32)
        self._write_blob("matches", image_pair_to_pair_id(image_id1, image_id2), m)

    def clear_matches(self) -> None:
        self._conn.execute("DELETE FROM matches")

    def write_two_view_geometry(self, image_id1: int, image_id2: int, geometry: TwoViewGeometry,
                                replace: bool = False) -> None:
        """database_sqlite.cc:1128-1190: stored for the (smaller id, larger id) order -- the geometry is
        inverted when the pair comes the other way round -- matrices as row-major float64 blobs, and no
        matrix blobs at all when there are no inlier matches."""
        g = geometry.inverted() if image_id1 > image_id2 else geometry
        m = np.ascontiguousarray(np.asarray(g.inlier_matches, dtype=np.uint32).reshape(-1, 2))
        blobs = [np.ascontiguousarray(np.asarray(a, dtype=np.float64)).tobytes() if m.shape[0] else b""
                 for a in (g.F, g.E, g.H, g.qvec, g.tvec)]
        verb = "INSERT OR REPLACE" if replace else "INSERT"
        self._conn.execute(f"{verb} INTO two_view_geometries(pair_id, rows, cols, data, config, F, E, H, qvec, tvec) "
                           "VALUES(?, ?, ?, ?, ?, ?, ?, ?, ?, ?)",
                           (image_pair_to_pair_id(image_id1, image_id2), int(m.shape[0]), 2, m.tobytes(),
                            int(g.config), *blobs))

    def clear_two_view_geometries(self) -> None:
        self._conn.execute("DELETE FROM two_view_geometries")

    # ------------------------------------------------------------------ reads
    def _count(self, table: str) -> int:
        return int(self._conn.execute(f"SELECT COUNT(*) FROM {table}").fetchone()[0])

    def num_cameras(self) -> int:
        return self._count("cameras")

    def num_images(self) -> int:
        return self._count("images")

    def num_matched_image_pairs(self) -> int:  # database_sqlite.cc:552-562 (empty rows count)
        return self._count("matches")

    def num_keypoints(self) -> int:
        r = self._conn.execute("SELECT SUM(rows) FROM keypoints").fetchone()[0]
        return int(r or 0)

    def num_descriptors(self) -> int:
        r = self._conn.execute("SELECT SUM(rows) FROM descriptors").fetchone()[0]
        return int(r or 0)

    def num_matches(self) -> int:
        r = self._conn.execute("SELECT SUM(rows) FROM matches").fetchone()[0]
        return int(r or 0)

    def _exists(self, table: str, keycol: str, key: int) -> bool:
        return self._conn.execute(f"SELECT 1 FROM {table} WHERE {keycol} = ?", (int(key),)).fetchone() is not None

    def exists_camera(self, camera_id: int) -> bool:
        return self._exists("cameras", "camera_id", camera_id)

    def exists_image(self, image_id: int) -> bool:
        return self._exists("images", "image_id", image_id)

    def exists_keypoints(self, image_id: int) -> bool:
        return self._exists("keypoints", "image_id", image_id)

    def exists_descriptors(self, image_id: int) -> bool:
        return self._exists("descriptors", "image_id", image_id)

    def exists_matches(self, image_id1: int, image_id2: int) -> bool:
        return self._exists("matches", "pair_id", image_pair_to_pair_id(image_id1, image_id2))

    def read_all_images(self) -> List[Image]:
        rows = self._conn.execute("SELECT image_id, name, camera_id FROM images ORDER BY image_id").fetchall()
        return [Image(name=n, camera_id=int(c), image_id=int(i)) for i, n, c in rows]

    def read_camera(self, camera_id: int) -> Camera:
        r = self._conn.execute("SELECT model, width, height, params, prior_focal_length FROM cameras "
                               "WHERE camera_id = ?", (int(camera_id),)).fetchone()
        if r is None:
            raise KeyError(camera_id)
        name = {v: k for k, v in CAMERA_MODEL_IDS.items()}[int(r[0])]
        return Camera(model=name, width=int(r[1]), height=int(r[2]),
                      params=np.frombuffer(r[3], dtype=np.float64).tolist(), camera_id=int(camera_id),
                      has_prior_focal_length=bool(r[4]))

    def _read_blob(self, table: str, keycol: str, key: int, dtype) -> np.ndarray:
        r = self._conn.execute(f"SELECT rows, cols, data FROM {table} WHERE {keycol} = ?", (int(key),)).fetchone()
        if r is None:
            return np.zeros((0, 0), dtype=dtype)
        rows, cols, data = r
        return np.frombuffer(data or b"", dtype=dtype).reshape(int(rows), int(cols)).copy()

    def read_keypoints(self, image_id: int) -> np.ndarray:
        return self._read_blob("keypoints", "image_id", image_id, np.float32)

    def read_descriptors(self, image_id: int) -> np.ndarray:
        return self._read_blob("descriptors", "image_id", image_id, np.uint8)

    def read_matches(self, image_id1: int, image_id2: int) -> np.ndarray:
        m = self._read_blob("matches", "pair_id", image_pair_to_pair_id(image_id1, image_id2), np.uint32)
        m = m.reshape(-1, 2)
        return m[:, ::-1].copy() if image_id1 > image_id2 else m

    @staticmethod
    def _geometry_from_row(rows, data, config, F, E, H, qvec, tvec) -> TwoViewGeometry:
        def mat(blob, shape):
            return np.frombuffer(blob, dtype=np.float64).reshape(shape).copy() if blob else np.zeros(shape)
        m = np.frombuffer(data or b"", dtype=np.uint32).reshape(int(rows), 2).copy() if rows else np.zeros((0, 2), np.uint32)
        q = mat(qvec, (4,)) if qvec else np.array([1.0, 0.0, 0.0, 0.0])
        return TwoViewGeometry(int(config), mat(F, (3, 3)), mat(E, (3, 3)), mat(H, (3, 3)), q, mat(tvec, (3,)), m)

    def exists_two_view_geometry(self, image_id1: int, image_id2: int) -> bool:
        return self._exists("two_view_geometries", "pair_id", image_pair_to_pair_id(image_id1, image_id2))

    def read_two_view_geometry(self, image_id1: int, image_id2: int) -> TwoViewGeometry:
        """database_sqlite.cc:839-881 (inverted back when asked for as (larger id, smaller id))."""
        r = self._conn.execute("SELECT rows, data, config, F, E, H, qvec, tvec FROM two_view_geometries "
                               "WHERE pair_id = ?", (image_pair_to_pair_id(image_id1, image_id2),)).fetchone()
        if r is None:
            return TwoViewGeometry()
        g = self._geometry_from_row(*r)
        return g.inverted() if image_id1 > image_id2 else g

    def read_two_view_geometries(self) -> List[Tuple[Tuple[int, int], TwoViewGeometry]]:
        """Every row in its stored (smaller id, larger id) orientation (database_sqlite.cc:883-930)."""
        out = []
        for pid, *r in self._conn.execute("SELECT pair_id, rows, data, config, F, E, H, qvec, tvec FROM "
                                          "two_view_geometries ORDER BY pair_id"):
            out.append((pair_id_to_image_pair(int(pid)), self._geometry_from_row(*r)))
        return out

    def num_verified_image_pairs(self) -> int:
        return self._count("two_view_geometries")

    def read_all_matches(self) -> List[Tuple[Tuple[int, int], np.ndarray]]:
        out = []
        for pid, rows, cols, data in self._conn.execute("SELECT pair_id, rows, cols, data FROM matches ORDER BY pair_id"):
            m = np.frombuffer(data or b"", dtype=np.uint32).reshape(int(rows), 2 if rows else 2)
            out.append((pair_id_to_image_pair(int(pid)), m.copy()))
        return out


def _database_cls():
    return _pycolmap.Database if _pycolmap is not None else Database


def _camera_cls():
    return _pycolmap.Camera if _pycolmap is not None else Camera


def _image_cls():
    return _pycolmap.Image if _pycolmap is not None else Image


def make_camera(model: str, width: int, height: int, params: Sequence[float]):
    """pycolmap.Camera(model=..., width=..., height=..., params=...) or the sqlite3 stand-in."""
    return _camera_cls()(model=model, width=width, height=height, params=list(params))


class ColmapDatabase:
    """Same surface as the reference wrapper (vit_colmap/database/colmap_db.py:6-75)."""

    def __init__(self, db_path: str) -> None:
        cls = _database_cls()
        try:
            self.db = cls.open(str(db_path))
        except TypeError:  # pycolmap 3.12 instance-method API (colmap_db.py:13-16)
            self.db = cls()
            self.db.open(str(db_path))

    def add_pinhole_camera(self, width: int, height: int, fx: float, fy: float, cx: float, cy: float) -> int:
        return self.db.write_camera(make_camera("PINHOLE", width, height, [fx, fy, cx, cy]))

    def add_image(self, name: str, camera_id: int) -> int:
        return self.db.write_image(_image_cls()(name=name, camera_id=camera_id))

    def add_keypoints(self, image_id: int, kpts: np.ndarray) -> None:
        self.db.write_keypoints(image_id, kpts.astype(np.float32))

    def add_descriptors(self, image_id: int, desc: np.ndarray) -> None:
        self.db.write_descriptors(image_id, desc.astype(np.uint8))

    def add_matches(self, image_id1: int, image_id2: int, pairs: np.ndarray) -> None:
        self.db.write_matches(image_id1, image_id2, pairs.astype(np.uint32))

    def commit(self) -> None:
        pass

    def close(self) -> None:
        """Extension: release the SQLite handle (the reference leaves this to garbage collection)."""
        if getattr(self, "db", None) is not None:
            self.db.close()
            self.db = None

    @staticmethod
    @contextmanager
    def open_database(db_path: str):
        cls = _database_cls()
        try:
            with cls.open(str(db_path)) as db:
                yield db
        except (TypeError, AttributeError):
            db = cls()
            db.open(str(db_path))
            try:
                yield db
            finally:
                db.close()

    @staticmethod
    def get_db_count(db, attr_name: str) -> int:
        attr = getattr(db, attr_name)
        return attr() if callable(attr) else attr
