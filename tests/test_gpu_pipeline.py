"""End-to-end tests through the reference-facing Pipeline / extractor API on the GPU.  The first one
is the reference's own tests/test_smoke_e2e.py restated against this package."""
import sqlite3

import numpy as np
import pytest

from test_boundary import _checkerboard, _write_shifted

pytestmark = pytest.mark.gpu


def test_smoke_e2e_dummy_extractor(tmp_path):
    """tests/test_smoke_e2e.py:20-76: dummy extractor, PINHOLE, matching on, reconstruction off."""
    from vit_colmap_b200.database import ColmapDatabase
    from vit_colmap_b200.pipeline.run_pipeline import Pipeline
    from vit_colmap_b200.utils.config import Config
    images = _write_shifted(tmp_path)
    db_path = tmp_path / "database.db"
    config = Config()
    config.extractor.extractor_type = "dummy"
    config.camera.model = "PINHOLE"
    config.do_matching = True
    config.do_reconstruction = False
    result = Pipeline(config).run(images, tmp_path / "out", db_path)
    assert db_path.exists() and result is None
    with ColmapDatabase.open_database(str(db_path)) as db:
        assert ColmapDatabase.get_db_count(db, "num_cameras") >= 1
        assert ColmapDatabase.get_db_count(db, "num_images") == 3
        assert ColmapDatabase.get_db_count(db, "num_matched_image_pairs") >= 1
        for image_id in (1, 2, 3):
            assert db.exists_keypoints(image_id) and db.exists_descriptors(image_id)
    # the fake backend is matched under COLMAP's L2 (index) rule: identical grid cells pair up one-to-one
    con = sqlite3.connect(str(db_path))
    rows = [r for (r,) in con.execute("SELECT rows FROM matches")]
    n_kp = con.execute("SELECT rows FROM keypoints WHERE image_id=1").fetchone()[0]
    assert len(rows) == 3 and all(r == n_kp for r in rows)


@pytest.fixture(scope="module")
def vit_s():
    from vit_colmap_b200.features.vit_extractor import ViTExtractor
    return ViTExtractor(model_name="dinov2_vits14", num_keypoints=1024, descriptor_dim=128, synthetic_seed=0)


def test_run_inference_contract(vit_s):
    """tests/test_vit_integration.py:18-60: dtypes, column counts, equal lengths on a 476x644 image."""
    rng = np.random.default_rng(0)
    img = rng.integers(0, 256, (480, 650, 3), dtype=np.uint8)   # resized to 476x644 on the host
    kps, desc = vit_s._run_inference(img)
    assert kps.dtype == np.float32 and desc.dtype == np.uint8
    assert kps.ndim == 2 and kps.shape[1] == 2 and desc.shape[1] == 128 and len(kps) == len(desc) > 0
    assert kps[:, 0].min() >= 0 and kps[:, 0].max() <= 650 and kps[:, 1].max() <= 480


def test_run_inference_matches_reference_postprocessing(vit_s):
    """The extractor's output equals the oracle post-processing applied to the extractor's own patch
    tokens (keypoints bit-exact apart from near-ties; descriptors within one quantisation step)."""
    import torch
    from oracle import extract_oracle as eo
    rng = np.random.default_rng(1)
    img = rng.integers(0, 256, (518, 518, 3), dtype=np.uint8)
    kps, desc = vit_s._run_inference(img)
    tokens = vit_s.model.forward_patch_tokens(torch.from_numpy(img).cuda()[None], bgr=True)[0].cpu()
    fm = tokens.reshape(37, 37, -1).permute(2, 0, 1).contiguous()
    ref = eo.dense_to_sparse(fm, (518, 518), (518, 518), 1024, 128, "harris",
                             projection=vit_s.descriptor_projection.cpu())
    a = {tuple(k) for k in kps.tolist()}
    b = {tuple(k) for k in ref["keypoints"].tolist()}
    assert len(a ^ b) <= 0.02 * len(b)          # near-ties of the score map (gap < 1e-5) may flip
    pos = {tuple(k): i for i, k in enumerate(ref["keypoints"].tolist())}
    common = [(i, pos[tuple(k)]) for i, k in enumerate(kps.tolist()) if tuple(k) in pos]
    gi, ri = zip(*common)
    diff = np.abs(desc[list(gi)].astype(int) - ref["desc_u8"][list(ri)].astype(int))
    assert diff.max() <= 1


def test_vit_pipeline_c1_checkerboards(tmp_path):
    """Config C1 shape: 8 translated 518x518 checkerboards, ViT-S/14, exhaustive matching into the DB."""
    import cv2
    from vit_colmap_b200.database import ColmapDatabase
    from vit_colmap_b200.features.vit_extractor import ViTExtractor
    from vit_colmap_b200.matching import match_exhaustive
    d = tmp_path / "images"
    d.mkdir()
    base = _checkerboard(518, 518, 37)
    for i in range(8):
        m = np.float32([[1, 0, 25 * i], [0, 1, 15 * i]])
        cv2.imwrite(str(d / f"img_{i:03d}.png"), cv2.warpAffine(base, m, (518, 518)))
    db_path = tmp_path / "db.db"
    ex = ViTExtractor(model_name="dinov2_vits14", num_keypoints=2048, descriptor_dim=128, synthetic_seed=0, batch_size=4)
    ex.extract(d, db_path, "SIMPLE_PINHOLE")
    res = match_exhaustive(str(db_path))
    assert len(res) == 28
    con = sqlite3.connect(str(db_path))
    assert con.execute("SELECT COUNT(*) FROM images").fetchone()[0] == 8
    assert con.execute("SELECT COUNT(*) FROM keypoints").fetchone()[0] == 8
    assert con.execute("SELECT COUNT(*) FROM matches").fetchone()[0] == 28
    rows = con.execute("SELECT rows, cols FROM descriptors").fetchall()
    assert all(c == 128 and r > 0 for r, c in rows)
    with ColmapDatabase.open_database(str(db_path)) as db:
        k = db.read_keypoints(1)
        assert k.dtype == np.float32 and k.shape[1] == 2


def test_trainable_vit_pipeline_into_database(tmp_path):
    """`trainable_vit` switch value end to end: 5 images (one unreadable, .ppm accepted), SIMPLE_RADIAL
    default camera (trainable_vit_extractor.py:302,323-341), 6-column keypoints and d-column uint8
    descriptors in the COLMAP tables, then exhaustive matching over them."""
    import cv2
    from vit_colmap_b200.database import ColmapDatabase
    from vit_colmap_b200.features.trainable_vit_extractor import TrainableViTExtractor
    from vit_colmap_b200.matching import match_exhaustive
    d = tmp_path / "images"
    d.mkdir()
    rng = np.random.default_rng(0)
    base = cv2.GaussianBlur(rng.integers(0, 256, (240, 300, 3), dtype=np.uint8), (0, 0), 2.0)
    for i in range(4):
        m = np.float32([[1, 0, 6 * i], [0, 1, 4 * i]])
        cv2.imwrite(str(d / f"img_{i:03d}.{'ppm' if i == 3 else 'png'}"), cv2.warpAffine(base, m, (300, 240)))
    (d / "img_broken.png").write_bytes(b"not an image")
    db_path = tmp_path / "db.db"
    ex = TrainableViTExtractor(model_name="dinov2_vits14", num_keypoints=512, descriptor_dim=64, score_threshold=0.1,
                               nms_radius=2, synthetic_seed=0, batch_size=3)
    ex.extract(d, db_path, "SIMPLE_RADIAL")
    con = sqlite3.connect(str(db_path))
    assert con.execute("SELECT COUNT(*) FROM images").fetchone()[0] == 4
    assert con.execute("SELECT model, params FROM cameras").fetchone()[0] == 2
    rows = con.execute("SELECT rows, cols FROM keypoints ORDER BY image_id").fetchall()
    assert len(rows) == 4 and all(c == 6 and 0 < r <= 512 for r, c in rows)
    drows = con.execute("SELECT rows, cols FROM descriptors ORDER BY image_id").fetchall()
    assert [r for r, _ in drows] == [r for r, _ in rows] and all(c == 64 for _, c in drows)
    with ColmapDatabase.open_database(str(db_path)) as db:
        k = db.read_keypoints(1)
        assert k.dtype == np.float32 and k.shape[1] == 6
        assert k[:, 0].min() >= 0 and k[:, 0].max() <= 299 and k[:, 1].max() <= 239
        assert np.all(np.diff(k[:, 4]) <= 0) and k[:, 4].min() > 0.1
    res = match_exhaustive(str(db_path))
    assert len(res) == 6
    assert con.execute("SELECT COUNT(*) FROM matches").fetchone()[0] == 6
    ex.backbone.close()


def test_hybrid_extractor_into_database(tmp_path):
    """Hybrid extractor (cv2 GFTT detector + ViT-S descriptors) end to end: _run_inference contract
    of hybrid_extractor.py:326-357, descriptors equal to the oracle chained on the same tokens, and
    rows in the COLMAP database."""
    import cv2
    import torch
    from oracle import extract_oracle as eo
    from vit_colmap_b200.database import ColmapDatabase
    from vit_colmap_b200.features.hybrid_extractor import ViTExtractor as Hybrid
    rng = np.random.default_rng(0)
    img_dir = tmp_path / "images"
    img_dir.mkdir()
    base = cv2.GaussianBlur(rng.integers(0, 256, (300, 400, 3), dtype=np.uint8), (0, 0), 2.0)
    base = cv2.normalize(base, None, 0, 255, cv2.NORM_MINMAX)
    for i in range(3):
        cv2.imwrite(str(img_dir / f"im{i}.png"), np.roll(base, 9 * i, axis=1))
    ex = Hybrid(model_name="dinov2_vits14", num_keypoints=300, descriptor_dim=64, detector_type="gftt",
                synthetic_seed=0)
    img = cv2.imread(str(img_dir / "im0.png"))
    kps, desc = ex._run_inference(img)
    assert kps.dtype == np.float32 and kps.shape[1] == 2 and desc.dtype == np.uint8
    assert desc.shape == (len(kps), 64) and len(kps) > 64
    # oracle on the same device tokens
    rimg, orig_wh, feat_wh = ex._resize(img)
    tok = ex.model.forward_patch_tokens(torch.from_numpy(rimg)[None].cuda(), bgr=True)
    gh, gw = rimg.shape[0] // 14, rimg.shape[1] // 14
    fm = tok[0].float().cpu().reshape(gh, gw, -1).permute(2, 0, 1).contiguous()
    ref = eo.hybrid_descriptors(fm, kps, orig_wh, feat_wh, 64, projection=ex.descriptor_projection.cpu())
    diff = np.abs(desc.astype(int) - ref["desc_u8"].astype(int))
    assert diff.max() <= 1 and (diff > 0).mean() < 5e-3
    db_path = tmp_path / "db.db"
    ex.extract(img_dir, db_path, "SIMPLE_PINHOLE")
    with ColmapDatabase.open_database(str(db_path)) as db:
        assert ColmapDatabase.get_db_count(db, "num_images") == 3
        for i in (1, 2, 3):
            assert db.exists_keypoints(i) and db.exists_descriptors(i)
            assert db.read_descriptors(i).shape[1] == 64
        k1 = db.read_keypoints(1)
    assert np.array_equal(k1, kps)


def test_pipelined_batches_equal_serial_batches(vit_s):
    """BatchPipeline.extract_batches (upload of batch k+1 under the kernels of batch k, two input
    slots) returns, in order, exactly what the serial _extract_batch returns; a failing batch is
    reported through on_error and does not disturb its neighbours."""
    rng = np.random.default_rng(3)
    sizes = [(224, 224), (224, 224), (154, 238), (224, 224), (224, 224)]
    batches = []
    for k, (h, w) in enumerate(sizes):
        imgs = [rng.integers(0, 256, (h, w, 3), dtype=np.uint8) for _ in range(1 + k % 3)]
        batches.append((imgs, (w, h), (w, h), None))
    vit_s._run_inference(batches[0][0][0])          # projection fitted before either schedule runs
    serial = [vit_s._extract_batch(b[0], b[1], b[2])[0] for b in batches]
    got = list(vit_s.extract_batches(batches))
    assert [g[0] for g in got] == list(range(len(batches)))
    for (_, out, _), ref in zip(got, serial):
        assert len(out) == len(ref)
        for (k1, d1), (k2, d2) in zip(out, ref):
            assert np.array_equal(k1, k2) and np.array_equal(d1, d2)
    # batch 1 is malformed (unequal image sizes): reported, skipped, the rest unaffected
    bad = list(batches)
    bad[1] = ([batches[1][0][0], batches[2][0][0]], (224, 224), (224, 224), None)
    errors = []
    got = list(vit_s.extract_batches(bad, on_error=lambda k, e: errors.append(k)))
    assert errors == [1] and got[1][1] is None
    for i in (0, 2, 3, 4):
        for (k1, d1), (k2, d2) in zip(got[i][1], serial[i]):
            assert np.array_equal(k1, k2) and np.array_equal(d1, d2)
    with pytest.raises(ValueError):
        list(vit_s.extract_batches(bad))


@pytest.mark.parametrize("rule", ["angular", "l2"])
def test_match_exhaustive_uploads_once_and_equals_oracle(tmp_path, rule):
    """match_exhaustive end to end over a database of SIFT-like uint8 descriptors (13 images, ragged
    counts, one image without descriptors): every written `matches` row equals the CPU oracle pair by
    pair (including the < min_num_matches blanking and the pair_id / column-swap rule), the resident
    path (one upload) and the per-pairing-block path (forced by max_resident_bytes) agree, a second
    call finds nothing to do (resume semantics), and the kernel-launch count shows the descriptors were
    converted ONCE rather than once per pair chunk."""
    from oracle import matching as om
    from vit_colmap_b200 import _lib
    from vit_colmap_b200.database import ColmapDatabase, Database
    from vit_colmap_b200.matching import FeatureMatchingOptions, match_exhaustive
    rng = np.random.default_rng(4)
    n_images, d = 13, 128
    counts = [int(c) for c in rng.integers(150, 420, n_images)]
    counts[5] = 0
    base = rng.standard_normal((600, d)).astype(np.float32)
    descs = []
    for n in counts:
        rows = np.abs(base[rng.permutation(600)[:n]] + 0.15 * rng.standard_normal((n, d)).astype(np.float32))
        descs.append(om.to_uint8_colmap(om.l2_normalize(rows)) if n else np.zeros((0, d), np.uint8))

    def make_db(path):
        db = ColmapDatabase(str(path))
        cam = db.add_pinhole_camera(640, 480, 500.0, 500.0, 320.0, 240.0)
        ids = [db.add_image(f"im{i:02d}.png", camera_id=cam) for i in range(n_images)]
        for i, image_id in enumerate(ids):
            if counts[i]:
                db.add_keypoints(image_id, rng.uniform(0, 400, (counts[i], 2)).astype(np.float32))
                db.add_descriptors(image_id, descs[i])
        db.close()
        return ids

    class Blocks:
        block_size = 5

    opts = FeatureMatchingOptions(max_ratio=0.85, max_distance=0.9, min_num_matches=15, rule=rule)
    lib = _lib.load()
    ids = make_db(tmp_path / "a.db")
    lib.vc_launch_count(1)
    res = match_exhaustive(str(tmp_path / "a.db"), opts, Blocks(), pairs_per_launch=16)
    launches_once = lib.vc_launch_count(1)
    make_db(tmp_path / "b.db")
    res_blocks = match_exhaustive(str(tmp_path / "b.db"), opts, Blocks(), max_resident_bytes=1)
    assert len(res) == n_images * (n_images - 1) // 2 and set(res) == set(res_blocks)
    n_nonempty = 0
    for (a, b), m in res.items():
        i, j = ids.index(a), ids.index(b)
        if rule == "l2":
            ref = om.match_l2_numpy(descs[i], descs[j], opts.max_ratio, opts.max_distance, True)
        else:
            ref = om.match_c(descs[i], descs[j], opts.max_ratio, opts.max_distance, True) if counts[i] and counts[j] \
                else np.zeros((0, 2), np.uint32)
        if len(ref) < 15:
            ref = np.zeros((0, 2), np.uint32)
        assert m.tolist() == ref.tolist(), (a, b)
        assert res_blocks[(a, b)].tolist() == m.tolist()
        n_nonempty += len(m) > 0
    assert n_nonempty > 30
    with Database.open(str(tmp_path / "a.db")) as db:
        assert db.num_matched_image_pairs() == len(res)
        for (a, b), m in list(res.items())[:20]:
            assert db.read_matches(a, b).tolist() == m.tolist()           # re-oriented to the (a, b) order asked for
            assert db.read_matches(b, a).tolist() == m[:, ::-1].tolist()  # stored once, columns follow min/max id
    assert match_exhaustive(str(tmp_path / "a.db"), opts, Blocks()) == {}
    # 78 pairs in chunks of <= 16 per pairing block: per chunk memset-free launches are the fused kernel,
    # its merge(s), emit, scan + pack; the u8 -> bf16 conversion (or the two l2 kernels) runs once
    from vit_colmap_b200.pairing import ExhaustivePairGenerator
    n_chunks = sum(-(-len(b) // 16) for b in ExhaustivePairGenerator(ids, 5).blocks())
    per_chunk = (launches_once - (2 if rule == "l2" else 1)) / n_chunks
    assert per_chunk == int(per_chunk) and 4 <= per_chunk <= 7, (launches_once, n_chunks)


def test_guided_matching_updates_two_view_geometries(tmp_path):
    """guided_matching over a database whose two_view_geometries rows were written by a verifier
    (here: synthetic homographies): inlier matches are replaced by the guided result (equal to the
    oracle), rows without a usable model are left alone, a pair left with fewer than min_num_inliers
    becomes an empty TwoViewGeometry (feature_matching_utils.cc:490-493)."""
    from oracle import matching as om
    from _scenes import two_view_scene as _two_view_scene
    from vit_colmap_b200.database import ColmapDatabase, Database, TwoViewGeometry
    from vit_colmap_b200.matching import FeatureMatchingOptions, guided_matching
    rng = np.random.default_rng(6)
    n, d = 250, 128
    desc, kps, config, G = _two_view_scene(rng, n, n, d, "H")
    db = ColmapDatabase(str(tmp_path / "g.db"))
    cam = db.add_pinhole_camera(640, 480, 500.0, 500.0, 320.0, 240.0)
    ids = [db.add_image(f"im{i}.png", camera_id=cam) for i in range(3)]
    for k, image_id in enumerate(ids):
        src = min(k, 1)
        kp6 = np.zeros((n, 6), np.float32)
        kp6[:, :2] = kps[src, :n]
        db.add_keypoints(image_id, kp6)                      # 6-column keypoints: only (x, y) is used
        db.add_descriptors(image_id, desc[src, :n])
    db.close()
    with Database.open(str(tmp_path / "g.db")) as raw:
        seed = np.array([[0, 0]], np.uint32)
        raw.write_two_view_geometry(ids[0], ids[1], TwoViewGeometry(config, H=G, inlier_matches=seed))
        raw.write_two_view_geometry(ids[0], ids[2], TwoViewGeometry(1, inlier_matches=seed))      # DEGENERATE: skipped
        # image 2 == image 3 but the stored model is wrong for them: (almost) nothing survives the gate
        raw.write_two_view_geometry(ids[1], ids[2], TwoViewGeometry(config, H=G, inlier_matches=seed))
    opts = FeatureMatchingOptions(max_ratio=0.9, max_distance=1.0)
    out = guided_matching(str(tmp_path / "g.db"), opts, max_error=2.0)
    ref = om.match_guided_numpy(desc[0, :n], desc[1, :n], kps[0], kps[1], config, None, G, 2.0, "angular", 0.9, 1.0)
    assert set(out) == {(ids[0], ids[1]), (ids[1], ids[2])}
    assert out[(ids[0], ids[1])].tolist() == ref.tolist() and len(ref) > 100
    assert len(out[(ids[1], ids[2])]) == 0
    with Database.open(str(tmp_path / "g.db")) as raw:
        g01 = raw.read_two_view_geometry(ids[0], ids[1])
        assert g01.config == config and g01.inlier_matches.tolist() == ref.tolist() and np.allclose(g01.H, G)
        assert raw.read_two_view_geometry(ids[1], ids[0]).inlier_matches.tolist() == ref[:, ::-1].tolist()
        g12 = raw.read_two_view_geometry(ids[1], ids[2])
        assert g12.config == 0 and len(g12.inlier_matches) == 0
        assert raw.read_two_view_geometry(ids[0], ids[2]).inlier_matches.tolist() == seed.tolist()


def _check_extraction_and_matches(ex, db_path, images, grid, num_keypoints, options, block_size=50):
    """Shared by the C2 / C3 shaped tests: database rows of `ex.extract` against the oracle
    post-processing of the extractor's own patch tokens, then match_exhaustive against the C oracle on
    the descriptors that were written."""
    import cv2
    import torch
    from oracle import extract_oracle as eo
    from oracle import matching as om
    from vit_colmap_b200.database import Database
    from vit_colmap_b200.matching import match_exhaustive
    gh, gw = grid
    with Database.open(str(db_path)) as db:
        ids = [im.image_id for im in db.read_all_images()]
        assert len(ids) == len(images)
        kps = [db.read_keypoints(i) for i in ids]
        descs = [db.read_descriptors(i) for i in ids]
    for k in (0, len(images) - 1):
        img = cv2.imread(str(images[k]))
        h, w = img.shape[:2]
        res, orig_wh, resized_wh = ex._resize(img)
        tok = ex.model.forward_patch_tokens(torch.from_numpy(res).cuda()[None], bgr=True)[0].cpu()
        fm = tok.reshape(gh, gw, -1).permute(2, 0, 1).contiguous()
        ref = eo.dense_to_sparse(fm, orig_wh, resized_wh, num_keypoints, ex.descriptor_dim, "harris",
                                 projection=ex.descriptor_projection.cpu())
        a = {tuple(p) for p in kps[k].tolist()}
        b = {tuple(p) for p in ref["keypoints"].tolist()}
        assert len(a ^ b) <= 0.02 * len(b), (k, len(a), len(b))    # score-map near-ties may flip a cell
        pos = {tuple(p): i for i, p in enumerate(ref["keypoints"].tolist())}
        common = [(i, pos[tuple(p)]) for i, p in enumerate(kps[k].tolist()) if tuple(p) in pos]
        gi, ri = zip(*common)
        assert np.abs(descs[k][list(gi)].astype(int) - ref["desc_u8"][list(ri)].astype(int)).max() <= 1
        assert kps[k][:, 0].max() <= w and kps[k][:, 1].max() <= h

    class Blocks:
        pass
    Blocks.block_size = block_size
    res = match_exhaustive(str(db_path), options, Blocks())
    for (a, b), m in res.items():
        ref = om.match_c(descs[ids.index(a)], descs[ids.index(b)], options.max_ratio, options.max_distance,
                         options.cross_check)
        if len(ref) < options.min_num_matches:
            ref = np.zeros((0, 2), np.uint32)
        assert m.tolist() == ref.tolist(), (a, b)
    # The reference's uint8 rows (negatives clipped, vit_extractor.py:250) have norm ~ 512 / sqrt(2), so even
    # identical rows sit at acos(0.5) > max_distance = 0.7 and the default thresholds accept nothing -- as
    # they would in COLMAP.  Mutual-nearest-neighbour thresholds (ratio and distance tests switched off,
    # config C1's matching mode) exercise long lists on the same descriptors.
    from vit_colmap_b200.matching import FeatureMatchingOptions, match_descriptor_sets
    loose = FeatureMatchingOptions(max_ratio=1.05, max_distance=3.2, cross_check=True)
    pairs = [(i, j) for i in range(len(ids)) for j in range(i + 1, len(ids))]
    got = match_descriptor_sets(descs, pairs, loose)
    total = 0
    for (i, j), m in zip(pairs, got):
        assert m.tolist() == om.match_c(descs[i], descs[j], 1.05, 3.2, True).tolist(), (i, j)
        total += len(m)
    assert total > 20 * len(pairs)
    return res, kps


def test_c2_hpatches_shaped_sequence(tmp_path):
    """BASELINE config C2 shape: one 6-image sequence (800x600 -> 798x588, grid 42 x 57), ViT-B/14 width,
    2048 keypoints, PCA 768 -> 128 fitted on the first image, ratio-test matching of the 15 pairs; images
    2..6 are homography warps of the first (corner jitter <= 15 %).  Depth-2 synthetic weights keep the
    test short; every stage still runs at C2's sizes."""
    import cv2
    from oracle import vit_oracle as vo
    from vit_colmap_b200.features.vit_extractor import ViTExtractor
    from vit_colmap_b200.matching import FeatureMatchingOptions
    rng = np.random.default_rng(116)
    d = tmp_path / "seq"
    d.mkdir()
    base = cv2.GaussianBlur(rng.integers(0, 256, (600, 800, 3), dtype=np.uint8), (0, 0), 3.0)
    base = cv2.normalize(base, None, 0, 255, cv2.NORM_MINMAX)
    corners = np.float32([[0, 0], [800, 0], [800, 600], [0, 600]])
    files = []
    for i in range(6):
        img = base
        if i:
            jitter = rng.uniform(-0.15, 0.15, (4, 2)).astype(np.float32) * np.float32([800, 600])
            img = cv2.warpPerspective(base, cv2.getPerspectiveTransform(corners, corners + jitter), (800, 600))
        # HPatches sequences hold 1.ppm .. 6.ppm, but the `vit` extractor only lists jpg / png / bmp / tiff
        # (vit_extractor.py:684-687; the trainable one adds .ppm) -- kept as is, so the sequence is written as png
        files.append(d / f"{i + 1}.png")
        cv2.imwrite(str(files[-1]), img)
    sd = vo.synthetic_state_dict("dinov2_vitb14", seed=2, depth=2)
    ex = ViTExtractor(model_name="dinov2_vitb14", num_keypoints=2048, descriptor_dim=128, state_dict=sd, batch_size=4)
    ex.extract(d, tmp_path / "c2.db", "SIMPLE_PINHOLE")
    opts = FeatureMatchingOptions(max_ratio=0.8, max_distance=0.7, cross_check=True, min_num_matches=15)
    res, kps = _check_extraction_and_matches(ex, tmp_path / "c2.db", files, (42, 57), 2048, opts)
    assert len(res) == 15 and all(150 < len(k) <= 2048 for k in kps)


def test_c3_dtu_shaped_images_into_database(tmp_path):
    """BASELINE config C3 shape: 1600x1200 images (resized to 1596x1190 on the host, grid 85 x 114,
    T = 9691 tokens), ViT-B/14 width, num_keypoints 8192 (the reference's binning + greedy NMS yields
    ~1.7k), exhaustive matching into the COLMAP database; 4 images, depth-2 synthetic weights."""
    import cv2
    from oracle import vit_oracle as vo
    from vit_colmap_b200.features.vit_extractor import ViTExtractor
    from vit_colmap_b200.matching import FeatureMatchingOptions
    rng = np.random.default_rng(49)
    d = tmp_path / "scan"
    d.mkdir()
    big = cv2.GaussianBlur(rng.integers(0, 256, (1300, 1800, 3), dtype=np.uint8), (0, 0), 2.5)
    big = cv2.normalize(big, None, 0, 255, cv2.NORM_MINMAX)
    yy, xx = np.mgrid[0:1300, 0:1800]
    big = np.where((((yy // 64) + (xx // 64)) % 2 == 0)[..., None], big, 255 - big).astype(np.uint8)   # noise + checker
    files = []
    for i in range(4):
        files.append(d / f"rect_{i + 1:03d}_3_r5000.png")     # DTU naming, scripts/run_DTU_vit.sh
        cv2.imwrite(str(files[-1]), big[14 * i:14 * i + 1200, 28 * i:28 * i + 1600])
    sd = vo.synthetic_state_dict("dinov2_vitb14", seed=3, depth=2)
    ex = ViTExtractor(model_name="dinov2_vitb14", num_keypoints=8192, descriptor_dim=128, state_dict=sd, batch_size=2)
    ex.extract(d, tmp_path / "c3.db", "SIMPLE_PINHOLE")
    con = sqlite3.connect(str(tmp_path / "c3.db"))
    model, w, h, params = con.execute("SELECT model, width, height, params FROM cameras").fetchone()
    assert (model, w, h) == (0, 1600, 1200) and np.frombuffer(params, np.float64).tolist() == [1600.0, 800.0, 600.0]
    opts = FeatureMatchingOptions(max_ratio=0.8, max_distance=0.7, cross_check=True, min_num_matches=15)
    res, kps = _check_extraction_and_matches(ex, tmp_path / "c3.db", files, (85, 114), 8192, opts)
    assert len(res) == 6 and all(1000 < len(k) <= 8192 for k in kps)
    assert con.execute("SELECT COUNT(*) FROM matches").fetchone()[0] == 6
