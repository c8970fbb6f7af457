"""libptscene.so (host scene library): bit-exact against the reference's own host functions where they are built
(oracle/_ref/libref_host.so, this container), against golden vectors generated from them everywhere else, plus
the library's own error handling."""
import ctypes as C
import os

import numpy as np
import pytest

from conftest import have_ref
from helpers import ROOT, from_bits, sha
from opengl_ray_tracing_b200 import scene as S, scenes

f32p = C.POINTER(C.c_float)


def fp(a):
    return a.ctypes.data_as(f32p)


@pytest.fixture(scope="module")
def ref():
    if not have_ref():
        pytest.skip("oracle/_ref not built (needs /root/reference)")
    L = C.CDLL(os.path.join(ROOT, "oracle", "_ref", "libref_host.so"))
    L.ref_build_bvh.restype = C.c_int
    L.ref_camera.argtypes = [C.c_float] * 3 + [f32p, f32p]
    return L


# ---------------- golden vectors (generated from the reference's sources by tests/golden/make_golden.py) ----------
def test_transform_matrix_golden(golden):
    for k in golden[0]["transform"]:
        out = S.getTransformMatrix(from_bits(k["rot"]), from_bits(k["trans"]), from_bits(k["scale"]))
        assert out.view(np.uint32).tolist() == k["bits"]


def test_camera_golden(golden):
    for k in golden[0]["camera"]:
        eye, cam = S.camera(k["up"], k["rot"], k["r"])
        assert eye.view(np.uint32).tolist() == k["eye"] and cam.view(np.uint32).tolist() == k["cam"]


def test_camera_default_is_reference_default():
    eye, cam = S.camera()                    # upAngle 0, rotateAngle 0, r 4 (IS/main.cpp:147-149)
    assert np.allclose(eye, [0, 0, 4]) and np.allclose(cam.reshape(4, 4)[3], [0, 0, 4, 1])


@pytest.mark.parametrize("threads", [1, 4])
def test_bvh_golden(golden, threads):
    for k in golden[0]["bvh"]:
        sc = S.Scene()
        scenes._populate(sc, "lambert", k["freq"])
        assert sc.nTriangles == k["nTriangles"]
        nn = sc.buildBVHwithSAH(k["leaf"], threads) if k["sah"] else sc.buildBVH(k["leaf"], threads)
        assert nn == k["nNodes"]
        assert sha(sc.nodes_encoded()) == k["nodes_sha256"], "node array differs from the reference builder's"
        assert sha(sc.triangles_encoded()) == k["tris_sha256"], "triangle order differs from the reference builder's"


def test_hdr_cache_golden(golden):
    for k in golden[0]["hdr_cache"]:
        hdr = S.synthHdr(k["w"], k["h"], k["seed"])
        assert sha(hdr) == k["hdr_sha256"], "procedural HDR generator is not deterministic"
        assert sha(S.calculateHdrCache(hdr, k["w"], k["h"])) == k["cache_sha256"]


# ---------------- live comparison with the reference's compiled host functions -------------------------------
def test_transform_and_camera_vs_ref(ref):
    rng = np.random.default_rng(0)
    for k in range(100):
        r = (rng.uniform(-180, 180, 3) if k > 3 else np.zeros(3)).astype(np.float32)
        t, s = rng.normal(size=3).astype(np.float32), rng.uniform(0.01, 20, 3).astype(np.float32)
        b = np.empty(16, np.float32)
        ref.ref_transform_matrix(fp(r), fp(t), fp(s), fp(b))
        assert S.getTransformMatrix(r, t, s).tobytes() == b.tobytes()
        up, rot, rr = rng.uniform(-89, 89), rng.uniform(-360, 360), rng.uniform(0.5, 10)
        e2, c2 = np.empty(3, np.float32), np.empty(16, np.float32)
        ref.ref_camera(up, rot, rr, fp(e2), fp(c2))
        e, c = S.camera(up, rot, rr)
        assert e.tobytes() == e2.tobytes() and c.tobytes() == c2.tobytes()


@pytest.mark.parametrize("smooth", [0, 1])
def test_read_obj_vs_ref(ref, tmp_path, smooth):
    path = str(tmp_path / "ico.obj")
    assert S.lib().pts_write_obj_icosphere(path.encode(), 9, 0.15, 4) == 0
    m = S.Material(baseColor=(0.2, 0.9, 0.4), roughness=0.3, metallic=0.5, clearcoat=1.0)
    tr = S.getTransformMatrix((10, 20, 30), (0.3, -1.6, 0.1), (1.5, 1.2, 0.7))
    sc = S.Scene()
    n = sc.readObj(path, m, tr, smooth)
    out = np.empty((n, 36), np.float32)
    ref.ref_read_obj.argtypes = [C.c_char_p, f32p, f32p, C.c_int, f32p, C.c_int]
    assert ref.ref_read_obj(path.encode(), fp(m.to_floats()), fp(tr), smooth, fp(out), n) == n == 20 * 81
    assert sc.triangles_encoded().tobytes() == out.tobytes()
    sc2 = S.Scene()                                      # the in-memory path is the same pipeline
    sc2.addIcosphere(9, 0.15, 4, m, tr, smooth)
    assert sc2.triangles_encoded().tobytes() == out.tobytes()


def test_obj_face_formats_vs_ref(ref, tmp_path):
    path = str(tmp_path / "f.obj")
    with open(path, "w") as f:
        f.write("# comment\nv 0 0 0\nv 1 0 0\nv 0 2 0\nv 0 0 3\nvt 0 0\nvn 0 0 1\n"
                "f 1 2 3\nf 1/1 2/1 4/1\nf 2/1/1 3/1/1 4/1/1\nf 1 3 4 2\n")
    m, tr = S.Material(), S.getTransformMatrix((0, 0, 0), (0, 0, 0), (1, 1, 1))
    sc = S.Scene()
    n = sc.readObj(path, m, tr, False)
    out = np.empty((n, 36), np.float32)
    ref.ref_read_obj.argtypes = [C.c_char_p, f32p, f32p, C.c_int, f32p, C.c_int]
    assert ref.ref_read_obj(path.encode(), fp(m.to_floats()), fp(tr), 0, fp(out), n) == n == 4
    assert sc.triangles_encoded().tobytes() == out.tobytes()


@pytest.mark.parametrize("sah", [0, 1])
def test_bvh_vs_ref_with_ties(ref, sah):
    """Quantised random soup: many equal centroids, so the result depends on std::sort's tie behaviour."""
    rng = np.random.default_rng(5)
    T = 4000
    tri = np.zeros((T, 36), np.float32)
    tri[:, :9] = np.round(rng.uniform(-2, 2, (T, 9)) * 4) / 4
    tri[:, 21:24] = 1
    tris = tri.copy()
    nodes = np.empty((2 * T + 2, 12), np.float32)
    nn = ref.ref_build_bvh(fp(tris), T, 8, sah, fp(nodes), len(nodes))
    for threads in (1, 4):
        sc = S.Scene()
        sc.addTriangles(tri)
        n1 = sc.buildBVHwithSAH(8, threads) if sah else sc.buildBVH(8, threads)
        assert n1 == nn and sc.nodes_encoded().tobytes() == nodes[:nn].tobytes()
        assert sc.triangles_encoded().tobytes() == tris.tobytes()


def test_hdr_loader_vs_ref(ref):
    ref.ref_hdr_load.argtypes = [C.c_char_p, C.POINTER(C.c_int), C.POINTER(C.c_int), f32p, C.c_long]
    base = os.environ.get("PT_REFERENCE_DIR", "/root/reference")
    for rel in ("DisneyBRDF/HDR/peppermint_powerplant_4k.hdr", "TestDemo/assets/HDR/white_cliff_top_1k.hdr"):
        p = os.path.join(base, rel)
        w, h, a = S.HDRLoader.load(p)
        w2, h2, b = C.c_int(), C.c_int(), np.empty((h, w, 3), np.float32)
        assert ref.ref_hdr_load(p.encode(), C.byref(w2), C.byref(h2), fp(b), b.size) == 0
        assert (w, h) == (w2.value, h2.value) and a.tobytes() == b.tobytes()
        c2 = np.empty_like(a)
        ref.ref_hdr_cache(fp(a), w, h, fp(c2))
        assert S.calculateHdrCache(a, w, h).tobytes() == c2.tobytes()


# ---------------- the library's own behaviour ------------------------------------------------------------------
def test_hdr_write_read_roundtrip(tmp_path):
    hdr = S.synthHdr(64, 32, 3)
    p = str(tmp_path / "t.hdr")
    S.writeHdr(p, hdr)
    w, h, back = S.HDRLoader.load(p)
    assert (w, h) == (64, 32)
    rel = np.abs(back - hdr).max(axis=2) / np.maximum(hdr.max(axis=2), 1e-6)
    assert rel.max() < 1.0 / 128.0                      # RGBE keeps 8 mantissa bits of the largest channel


def test_errors_do_not_exit(tmp_path):
    sc = S.Scene()
    with pytest.raises(S.SceneError, match="open failed"):
        sc.readObj(str(tmp_path / "missing.obj"), S.Material(), np.eye(4, dtype=np.float32).ravel(), True)
    bad = tmp_path / "bad.obj"
    bad.write_text("v 0 0 0\nv 1 0 0\nv 0 1 0\nf 1//1 2//1 3//1\n")
    with pytest.raises(S.SceneError, match="unsupported face"):
        sc.readObj(str(bad), S.Material(), np.eye(4, dtype=np.float32).ravel(), True)
    oob = tmp_path / "oob.obj"
    oob.write_text("v 0 0 0\nv 1 0 0\nv 0 1 0\nf 1 2 9\n")
    with pytest.raises(S.SceneError, match="out of range"):
        sc.readObj(str(oob), S.Material(), np.eye(4, dtype=np.float32).ravel(), True)
    with pytest.raises(S.SceneError):
        S.HDRLoader.load(str(bad))
    assert sc.nTriangles == 0 and sc.nNodes == 0


def test_scene_cache_roundtrip(tmp_path):
    sc = S.Scene()
    scenes._populate(sc, "is", 6)
    sc.buildBVHwithSAH(8)
    t, n = sc.triangles_encoded(), sc.nodes_encoded()
    p = str(tmp_path / "s.ptsc")
    sc.save(p)
    sc2 = S.Scene()
    sc2.load(p)
    assert sc2.triangles_encoded().tobytes() == t.tobytes() and sc2.nodes_encoded().tobytes() == n.tobytes()
    with open(p, "r+b") as f:
        f.truncate(100)
    with pytest.raises(S.SceneError):
        S.Scene().load(p)


def test_tree_structure_invariants():
    """Node 0 is the dummy (IS/main.cpp:780-786), root is node 1, leaves partition [0, T) in order, left = id + 1."""
    sc = S.Scene()
    scenes._populate(sc, "lambert", 10)
    nn = sc.buildBVHwithSAH(8, 2)
    nodes = sc.nodes_encoded()
    assert nodes[0, :6].tolist() == [255, 128, 0, 30, 0, 0] and nodes[0, 6:].tolist() == [1, 1, 0, 0, 1, 0]
    inner = nodes[1:, 3] == 0
    ids = np.arange(1, nn)
    assert (nodes[1:, 0][inner] == ids[inner] + 1).all()
    leaves = nodes[1:][~inner]
    assert (leaves[:, 3] <= 8).all() and (leaves[:, 3] >= 1).all()
    order = np.argsort(leaves[:, 4])
    first, cnt = leaves[order, 4].astype(int), leaves[order, 3].astype(int)
    assert first[0] == 0 and (first[1:] == first[:-1] + cnt[:-1]).all() and first[-1] + cnt[-1] == sc.nTriangles
    tris = sc.triangles_encoded()
    for leaf in leaves[:50]:                              # leaf boxes are the exact bounds of their triangles
        p = tris[int(leaf[4]):int(leaf[4] + leaf[3]), :9].reshape(-1, 3)
        assert (p.min(axis=0) == leaf[6:9]).all() and (p.max(axis=0) == leaf[9:12]).all()


def test_mesh_generator_counts_and_determinism():
    for n in (1, 3, 8):
        a, b = S.Scene(), S.Scene()
        tr = np.eye(4, dtype=np.float32).ravel()
        assert a.addIcosphere(n, 0.1, 9, S.Material(), tr) == 20 * n * n
        b.addIcosphere(n, 0.1, 9, S.Material(), tr)
        assert a.triangles_encoded().tobytes() == b.triangles_encoded().tobytes()
    # welded: every edge is shared by exactly two triangles (closed surface)
    a = S.Scene(); a.addIcosphere(4, 0.1, 9, S.Material(), np.eye(4, dtype=np.float32).ravel(), False)
    p = a.triangles_encoded()[:, :9].reshape(-1, 3, 3)
    edges = {}
    for t in p:
        for i in range(3):
            e = tuple(sorted([t[i].tobytes(), t[(i + 1) % 3].tobytes()]))
            edges[e] = edges.get(e, 0) + 1
    assert set(edges.values()) == {2}
