"""OBJ ingest: include/visionaray_b200/obj_loader.h follows load_obj() of the reference
(src/common/obj_loader.cpp:299-533, grammar src/common/obj_grammar.cpp:33-60).  The reference's loader needs Boost
(Spirit, Iostreams, Filesystem), absent here, so it cannot produce fixtures: the arrays are pinned against a
restatement of its rules below (fans, index remapping, degenerate rejection, geom_id = materials - 1, materials from
the mtl entries, padding), bit for bit; the frames rendered from the loaded arrays are pinned against the oracle fed
the same arrays (GPU tier).  Fixtures: tests/golden/obj/scene.obj (+ scene.mtl, checker.ppm; hand-written: quads, fans
with negative indices, CRLF, tabs, a degenerate face, an unknown material, a missing texture) and smooth.obj (a UV
sphere with vt / vn on every face)."""
import os

import numpy as np
import pytest

from oracle import bindings as O
from visionaray_b200 import scenes as S

OBJ = os.path.join(os.path.dirname(__file__), "golden", "obj")
f32 = np.float32


def _restate(path):
    """The rules of obj_loader.cpp, independently: returns (tris[n,3,3] as v1,e1,e2, geom_ids, shading normals, tex coords,
    materials [(ka, kd, ks, ns)], texture names)."""
    def words(line):
        line = line.rstrip("\r\n").strip(" \t")
        k = line.split(None, 1)
        return (k[0], k[1].strip(" \t") if len(k) > 1 else "") if k else ("", "")

    matlib, cur = {}, None
    verts, vts, vns = [], [], []
    tris, gids, sn, tc, mats, texnames = [], [], [], [], [], []
    geom_id = 0
    d = os.path.dirname(path)
    for line in open(path, newline=""):
        for line in line.split("\n"):
            key, rest = words(line)
            if not key or key.startswith("#"):
                continue
            if key == "mtllib":
                cur = None
                for ml in open(os.path.join(d, rest)):
                    k, r = words(ml)
                    if k == "newmtl":
                        cur = matlib.setdefault(r, dict(ka=(0.2, 0.2, 0.2), kd=(0.8, 0.8, 0.8), ks=(0.1, 0.1, 0.1), ns=32.0, map_kd=""))
                    elif cur is not None and k in ("Ka", "Kd", "Ks"):
                        cur[k.lower()] = tuple(float(x) for x in r.split()[:3])
                    elif cur is not None and k == "Ns":
                        cur["ns"] = float(r.split()[0])
                    elif cur is not None and k == "map_Kd":
                        cur["map_kd"] = r
            elif key == "usemtl":
                if rest in matlib:
                    m = matlib[rest]
                    mats.append((m["ka"], m["kd"], m["ks"], m["ns"]))
                    texnames.append(m["map_kd"])
                geom_id = len(mats) - 1 if mats else 0
            elif key == "v":
                verts.append(np.array([f32(x) for x in rest.split()[:3]], dtype=f32))
            elif key == "vt":
                vts.append(np.array([f32(x) for x in rest.split()[:2]], dtype=f32))
            elif key == "vn":
                vns.append(np.array([f32(x) for x in rest.split()[:3]], dtype=f32))
            elif key == "f":
                corners = []
                for tok in rest.split():
                    p = tok.split("/")
                    corners.append((int(p[0]), int(p[1]) if len(p) > 1 and p[1] else None, int(p[2]) if len(p) > 2 and p[2] else None))
                remap = lambda i, n: n + i if i < 0 else i - 1
                for last in range(2, len(corners)):
                    c = (corners[0], corners[last - 1], corners[last])
                    v1, v2, v3 = (verts[remap(x[0], len(verts))] for x in c)
                    e1, e2 = (v2 - v1).astype(f32), (v3 - v1).astype(f32)
                    cr = np.array([f32(e1[1] * e2[2]) - f32(e1[2] * e2[1]), f32(e1[2] * e2[0]) - f32(e1[0] * e2[2]),
                                   f32(e1[0] * e2[1]) - f32(e1[1] * e2[0])], dtype=f32)
                    if float(np.sqrt(f32(f32(f32(cr[0] * cr[0]) + f32(cr[1] * cr[1])) + f32(cr[2] * cr[2])))) == 0.0:
                        continue
                    tris.append((v1, e1, e2)); gids.append(len(mats) - 1 if mats else 0)
                    if all(x[1] is not None for x in c):
                        tc += [vts[remap(x[1], len(vts))] for x in c]
                    if all(x[2] is not None for x in c):
                        sn += [vns[remap(x[2], len(vns))] for x in c]
    while len(mats) <= geom_id:
        mats.append(((0.2, 0.2, 0.2), (0.8, 0.8, 0.8), (0.1, 0.1, 0.1), 32.0)); texnames.append("")
    return np.array(tris, dtype=f32), np.array(gids), np.array(sn, dtype=f32).reshape(-1, 3), np.array(tc, dtype=f32).reshape(-1, 2), mats, texnames


@pytest.mark.parametrize("name", ["scene.obj", "smooth.obj"])
def test_arrays_follow_the_reference_loader_rules(name):
    m = S.load_obj(os.path.join(OBJ, name))
    tris, gids, sn, tc, mats, texnames = _restate(os.path.join(OBJ, name))
    n = tris.shape[0]
    assert m.prims.shape[0] == n
    ids = m.prims.view(np.uint32)
    assert (ids[:, 1] == np.arange(n)).all() and (ids[:, 0] == gids).all()                  # prim_id = position, geom_id = material
    assert m.prims[:, 4:7].tobytes() == tris[:, 0].tobytes()                                  # v1, e1, e2 bit for bit
    assert m.prims[:, 8:11].tobytes() == tris[:, 1].tobytes() and m.prims[:, 12:15].tobytes() == tris[:, 2].tobytes()
    assert m.shading_normals[:, :3].tobytes() == sn.tobytes()
    assert m.tex_coords.shape[0] == 3 * n                                                     # padded to 3 per triangle
    assert m.tex_coords[:tc.shape[0]].tobytes() == tc.tobytes() and not m.tex_coords[tc.shape[0]:].any()
    assert m.materials.shape[0] == len(mats) and len(m.textures) == len(mats)
    for got, (ka, kd, ks, ns) in zip(m.materials, mats):
        assert np.array_equal(got["ca"], np.array(ka, dtype=f32)) and np.array_equal(got["cd"], np.array(kd, dtype=f32))
        assert np.array_equal(got["cs"], np.array(ks, dtype=f32)) and got["spec_exp"] == f32(ns)
        assert got["ka"] == 1.0 and got["kd"] == 1.0 and got["ks"] == 1.0 and got["kind"] == 0
    e1, e2 = m.prims[:, 8:11].astype(np.float64), m.prims[:, 12:15].astype(np.float64)
    g = np.cross(e1, e2); g /= np.linalg.norm(g, axis=1)[:, None]
    assert np.abs(m.geometric_normals[:, :3] - g).max() < 1e-6
    v = np.concatenate([m.prims[:, 4:7], m.prims[:, 4:7] + m.prims[:, 8:11], m.prims[:, 4:7] + m.prims[:, 12:15]])
    assert np.array_equal(m.bbox[0, :3], v.min(axis=0)) and np.array_equal(m.bbox[1, :3], v.max(axis=0))
    for t, tname in zip(m.textures, texnames):
        assert (t is not None) == (tname == "checker.ppm")


def test_fixture_specifics_and_errors():
    m = S.load_obj(os.path.join(OBJ, "scene.obj"))
    assert m.prims.shape[0] == 10                      # 2 (quad) + 4 (pyramid) + 3 (pentagon fan) + 1; the degenerate face is rejected
    assert m.prims.view(np.uint32)[:, 0].tolist() == [0, 0, 1, 1, 1, 1, 1, 1, 1, 2]       # the unknown material adds nothing
    assert m.shading_normals.shape[0] == 18            # only faces with normals on all three corners contribute (reference quirk)
    t = m.textures[0]
    assert t.data.shape == (6, 8, 4) and t.filter == 1 and tuple(t.address) == (0, 0)       # Linear, Wrap (obj_loader.cpp:392-393)
    assert t.data[2, 3].tolist() == [255, 0, 0, 255] and (t.data[..., 3] == 255).all()       # RGB8 -> RGBA8, alpha 1
    assert any("does_not_exist" in w for w in m.warnings) and any("degenerate" in w for w in m.warnings)
    assert any("missing.png" in w for w in m.warnings)
    assert m.materials["spec_exp"].tolist() == [8.0, 48.5, 32.0]
    with pytest.raises(RuntimeError):
        S.load_obj(os.path.join(OBJ, "no_such_file.obj"))


def _frame_case(name, kernel):
    m = S.load_obj(os.path.join(OBJ, name))
    sc = m.host_scene()
    sd = m.surface()
    eye = (1.6, 2.1, 2.9)
    fp = S.frame_params(kernel, 224, 160, eye, (0.0, 0.6, 0.0), num_bounces=3, n_materials=len(m.materials),
                        ambient=(1, 1, 1, 1) if kernel == 3 else (0.2, 0.2, 0.2, 1.0), pixel_sampler=2 if kernel == 3 else 0)
    return m, sc, sd, fp, S.point_light((2.5, 4.0, 3.0))


@pytest.mark.parametrize("name", ["scene.obj", "smooth.obj"])
def test_oracle_renders_the_loaded_model(name):
    m, sc, sd, fp, light = _frame_case(name, 1)
    img = O.oracle_frame_surface(sc, fp, sd, m.materials, light)
    assert np.isfinite(img).all() and 0.05 < float((img[..., :3] != np.array(fp.bg[:3], dtype=f32)).any(axis=2).mean()) < 0.95
    if name == "scene.obj":
        assert sd.normal_binding == 0                  # normals on some faces only: per-face geometric normals
    else:
        assert sd.normal_binding == 1 and sd.normals.shape[0] == 3 * sc.n_prims


@pytest.mark.gpu
@pytest.mark.parametrize("kernel", [0, 1, 3])
@pytest.mark.parametrize("name", ["scene.obj", "smooth.obj"])
def test_gpu_frames_of_the_loaded_model_match_the_oracle(gpu, name, kernel):
    m, sc, sd, fp, light = _frame_case(name, kernel)
    dev = gpu.Scene(sc)
    gpu.set_surface(dev, sd)
    want = O.oracle_frame_surface(sc, fp, sd, m.materials, light)
    for trav in (gpu.TRAVERSE_EXACT, gpu.TRAVERSE_WIDE):
        img = np.zeros((160, 224, 4), dtype=np.float32)
        gpu.frame(dev, fp, None, m.materials, light, img, traverse=trav)
        bad = float((np.abs(img - want).max(axis=2) > 1.0 / 255.0).mean())
        assert bad <= (0.0 if trav == gpu.TRAVERSE_EXACT else 3e-3), (name, kernel, trav, bad)
    dev.close()
    built = gpu.Scene.build(0, m.prims, None)          # the same model through the GPU builder
    gpu.set_surface(built, sd)
    img = np.zeros((160, 224, 4), dtype=np.float32)
    gpu.frame(built, fp, None, m.materials, light, img)
    assert float((np.abs(img - want).max(axis=2) > 1.0 / 255.0).mean()) <= 3e-3
    built.close()
