"""The .scene / OBJ / MTL loader (csrc/ptb_scene_loader.cpp, ptb_convert_scene_file / ptb_load_scene_file)
against the reference's own loader: loadScene + loadObject (tinyobjloader 2.0.0 underneath) as compiled into
oracle/_ref/ptref, whose `dump-scene` writes what getTris() / getObjects() / getMaterials() hold.  The bar is byte equality
of the .ptbs container — vertex transforms, normals, flipped v coordinates, material ids, object ranges, Heron surface
areas, emission normalisation, scene-file overrides, names.  tests/golden/cornel.ptbs is such a dump (committed), the
other comparisons run where the reference is compiled (this container), on OBJ text written by the test itself."""
import os

import numpy as np
import pytest

from oracle_lib import have_ref, run_ptref

REF_SHOWCASE = "/root/reference/showcase"
GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "cornel.ptbs")


def _convert(ptb, scene, out, base="", **kw):
    ptb.convert_scene_file(scene, out, base_dir=base, **kw)
    with open(out, "rb") as f:
        return f.read()


@pytest.fixture(scope="module")
def ptb():
    import ptb_b200
    return ptb_b200


@pytest.mark.skipif(not os.path.isfile(os.path.join(REF_SHOWCASE, "scenes", "cornel.scene")), reason="the reference's showcase/ assets are not on this machine")
def test_cornell_scene_equals_the_reference_loaders_dump(ptb, tmp_path):
    mine = _convert(ptb, os.path.join(REF_SHOWCASE, "scenes", "cornel.scene"), str(tmp_path / "c.ptbs"), REF_SHOWCASE)
    with open(GOLDEN, "rb") as f:
        assert mine == f.read()


OBJ = """# written by tests/test_scene_loader.py
mtllib m.mtl
o first object
v 0 0 0
v 1.5 0 0
v 1.5 2.25 0
v 0 2.25 1e-1
v -.5 -0.25 3
v 2.5E+0 -1.75e-2 3.125
v 0.333333343 0.1 0.7
v 12345.678 -0.000012 7
vn 0 0 1
vn 0 1 0
vn 0.577350 0.577350 -0.577350
vt 0 0
vt 1 0
vt 1 1
vt 0.25 0.75
usemtl red
f 1/1/1 2/2/1 3/3/1
f 1/1/1 2/2/1 3/3/1 4/4/2
usemtl lamp
f -4/1/3 -3/2/3 -2/3/3
g second group of faces
usemtl red
f 5/4/2 6/3/2 7/2/2 8/1/2
f 8/1/1 7/2/2 6/3/3
o third
f 2/2/1 4/4/1 6/1/1\r
usemtl nowhere
f 1/1/1 3/3/1 5/2/2
"""

MTL = """newmtl red
Ns 250.000000
Kd 0.800204 0.000000 0.458439
Ke 0 0 0
Ni 1.5

newmtl lamp
Kd 0.1 0.2 0.3
Ke 4.5 9.9 0.99
illum 2
newmtl dim
Ke 0.25 0.5 0.125
"""

SCENE = """[Models]
file: objs/a.obj
pos:  0.5,-250,3
scale: 300,2.5,-1
file: objs/a.obj
scale: 1,1,1

[Materials]
name: red
color: 1.0,0.25,0.8
roughness: 0.1
weights: 0.25,0.25,0.5
name: dim
emission: 3

[Camera]
pos: 1,2,-500
forward: 0.1,0.0,1
focus: 1.5
dof: 0.25
"""


def _write_tree(d):
    os.makedirs(os.path.join(d, "objs"))
    with open(os.path.join(d, "objs", "a.obj"), "w", newline="") as f:
        f.write(OBJ)
    with open(os.path.join(d, "objs", "m.mtl"), "w") as f:
        f.write(MTL)
    with open(os.path.join(d, "s.scene"), "w") as f:
        f.write(SCENE)
    return os.path.join(d, "s.scene")


def test_synthetic_scene_arrays(ptb, tmp_path):
    """What the text above must turn into, worked out by hand for the parts that need no reference: triangle count (two
    quads -> four triangles), the shorter-diagonal rule, relative indices, per-face material ids with the fallback for an
    unknown usemtl, shapes split at o / g, two instances of the model with their own fallback material each."""
    scene = _write_tree(str(tmp_path))
    _convert(ptb, scene, str(tmp_path / "s.ptbs"), str(tmp_path))
    s = ptb.scenes.read_ptbs(str(tmp_path / "s.ptbs"))
    assert s["pos"].shape == (18, 9) and len(s["objects"]) == 6 and len(s["materials"]) == 8
    names = [m["name"] for m in s["materials"]]
    assert names == ["None", "red", "lamp", "dim", "None", "red", "lamp", "dim"]
    assert [o["name"] for o in s["objects"]] == ["first object", "second group of faces", "third"] * 2
    assert [(o["start"], o["end"]) for o in s["objects"]] == [(0, 4), (4, 7), (7, 9), (9, 13), (13, 16), (16, 18)]
    # first model: materials 1.. ; unknown usemtl -> the shape's first material id is >= 0, so the face takes the shape's
    assert s["mat"].tolist() == [1, 1, 1, 2, 1, 1, 1, 1, 1] + [5, 5, 5, 6, 5, 5, 5, 5, 5]
    assert [o["materials"] for o in s["objects"]] == [[1, 2], [1], [1], [5, 6], [5], [5]]
    # transform of the first instance: v * scale + pos in fp32, two roundings
    v = np.array([1.5, 2.25, 0], np.float32) * np.array([300, 2.5, -1], np.float32) + np.array([0.5, -250, 3], np.float32)
    assert (s["pos"][0, 6:9].view(np.uint32) == v.view(np.uint32)).all()
    # quad 1-2-3-4: |v3-v1|^2 = 7.3125 < |v4-v2|^2 = 7.3225 -> triangles (1,2,3) (1,3,4)
    second = s["pos"][9:]  # the second instance: scale 1, and the position carried over from the first model
    v3 = np.array([1.5 + 0.5, 2.25 - 250, 0 + 3], np.float32)
    assert (second[1, 6:9] == v3).all() and (second[2, 3:6] == v3).all()
    # flipped v, emission normalised by its largest component, override by name on both instances
    assert s["uv"][0].tolist() == [0, 1, 1, 1, 1, 0]
    lamp, red, dim = s["materials"][2], s["materials"][5], s["materials"][7]
    assert lamp["emission"] == np.float32(9.9) and np.allclose(lamp["albedo"], [4.5 / 9.9, 1, 0.99 / 9.9])
    assert red["albedo"] == [1.0, 0.25, np.float32(0.8)] and red["weights"] == [0.25, 0.25, 0.5] and red["roughness"] == np.float32(0.1)
    assert dim["emission"] == 3.0 and dim["roughness"] == np.float32(0.1) and dim["weights"] == [0.25, 0.25, 0.5]  # fields carry over
    assert s["camera"] == {"pos": [1.0, 2.0, -500.0], "forward": [np.float32(0.1), 0.0, 1.0], "focus": 1.5, "dof": 0.25}


@pytest.mark.skipif(not have_ref(), reason="oracle/_ref/ptref not built (needs the reference sources)")
def test_synthetic_scene_equals_the_compiled_reference(ptb, tmp_path):
    scene = _write_tree(str(tmp_path))
    mine = _convert(ptb, scene, str(tmp_path / "mine.ptbs"), str(tmp_path))
    run_ptref(["--scene-file", scene, "--base", str(tmp_path), "dump-scene", str(tmp_path / "ref.ptbs")])
    with open(tmp_path / "ref.ptbs", "rb") as f:
        ref = f.read()
    assert mine == ref


@pytest.mark.parametrize("chunks", [2, 3, 7, 64])
def test_chunked_parse_equals_the_single_chunk_parse(ptb, tmp_path, monkeypatch, chunks):
    """Large files are cut into chunks of whole lines parsed by one thread each (relative indices resolved from prefix
    counts, usemtl / g / o replayed in file order); forcing that path on the small scene puts the cuts anywhere."""
    scene = _write_tree(str(tmp_path))
    monkeypatch.setenv("PTB_LOADER_THREADS", "1")
    one = _convert(ptb, scene, str(tmp_path / "one.ptbs"), str(tmp_path))
    monkeypatch.setenv("PTB_LOADER_THREADS", str(chunks))
    many = _convert(ptb, scene, str(tmp_path / "many.ptbs"), str(tmp_path))
    assert one == many


@pytest.mark.skipif(not (have_ref() and os.path.isfile(os.path.join(REF_SHOWCASE, "objects", "test.obj"))), reason="needs ptref and the reference's showcase/objects/test.obj")
def test_blender_export_equals_the_compiled_reference(ptb, tmp_path):
    """showcase/objects/test.obj: 1,932 triangles of a Blender export with its own .mtl (naming a texture file that is
    not shipped: both loaders go on without it), moved and scaled anisotropically"""
    scene = str(tmp_path / "t.scene")
    with open(scene, "w") as f:
        f.write("[Models]\nfile: objects/test.obj\npos: 3.25,-17,0.001\nscale: 30,31.5,-29\n\n[Camera]\npos: 0,0,-500\nforward: 0,0,1\nfocus: 1\n")
    mine = _convert(ptb, scene, str(tmp_path / "mine.ptbs"), REF_SHOWCASE)
    os.environ["PTB_LOADER_THREADS"] = "5"
    try:
        chunked = _convert(ptb, scene, str(tmp_path / "mine5.ptbs"), REF_SHOWCASE)
    finally:
        del os.environ["PTB_LOADER_THREADS"]
    run_ptref(["--scene-file", scene, "--base", REF_SHOWCASE, "dump-scene", str(tmp_path / "ref.ptbs")])
    with open(tmp_path / "ref.ptbs", "rb") as f:
        ref = f.read()
    assert mine == ref and chunked == ref


def test_refusals(ptb, tmp_path):
    d = str(tmp_path)
    os.makedirs(os.path.join(d, "objs"))
    with open(os.path.join(d, "objs", "m.mtl"), "w") as f:
        f.write("newmtl t\nKd 1 1 1\nmap_Kd -bm 0.5 wall tex.png\n")
    with open(os.path.join(d, "objs", "a.obj"), "w") as f:
        f.write("mtllib m.mtl\nv 0 0 0\nv 1 0 0\nv 1 1 0\nv 0 1 0\nv -1 0.5 0\nvn 0 0 1\nvt 0 0\nusemtl t\nf 1/1/1 2/1/1 3/1/1\n")
    with open(os.path.join(d, "s.scene"), "w") as f:
        f.write("[Models]\nfile: objs/a.obj\n")
    # a texture file that is not there: "could not load", untextured material, like loadTexture (texture.cpp:36-39)
    ptb.convert_scene_file(os.path.join(d, "s.scene"), os.path.join(d, "o.ptbs"), base_dir=d)
    with open(os.path.join(d, "objs", "wall tex.png"), "wb") as f:
        f.write(b"\x89PNG")
    with pytest.raises(ptb.PtbError, match="wall tex.png"):  # the file name keeps its blank; options are skipped
        ptb.convert_scene_file(os.path.join(d, "s.scene"), os.path.join(d, "o.ptbs"), base_dir=d)
    ptb.convert_scene_file(os.path.join(d, "s.scene"), os.path.join(d, "o.ptbs"), base_dir=d, ignore_textures=True)
    assert ptb.scenes.read_ptbs(os.path.join(d, "o.ptbs"))["pos"].shape == (1, 9)
    with pytest.raises(ptb.PtbError, match="cannot open"):
        ptb.convert_scene_file(os.path.join(d, "missing.scene"), os.path.join(d, "o.ptbs"))
    with open(os.path.join(d, "s2.scene"), "w") as f:
        f.write("[Models]\nfile: objs/none.obj\n")
    with pytest.raises(ptb.PtbError, match="Cannot open file"):
        ptb.convert_scene_file(os.path.join(d, "s2.scene"), os.path.join(d, "o.ptbs"), base_dir=d)


# ---- PNG textures (csrc/ptb_png.cpp) against stb_image as compiled into the reference -----------------------------------------
def _png_bytes(samples, colour, depth, filters, plte=None, trns=None):
    """samples: (H, W, channels) integer array of raw sample values; one filter type per row, cycling through `filters`"""
    import struct
    import zlib
    h, w, ch = samples.shape
    rows = []
    for y in range(h):
        s = samples[y].reshape(-1)
        if depth == 8:
            raw = bytes(int(v) for v in s)
        elif depth == 16:
            raw = b"".join(struct.pack(">H", int(v)) for v in s)
        else:
            bits = "".join(format(int(v), "0%db" % depth) for v in s)
            bits += "0" * (-len(bits) % 8)
            raw = bytes(int(bits[i:i + 8], 2) for i in range(0, len(bits), 8))
        rows.append(raw)
    bpp = max(1, ch * depth // 8)
    out = b""
    prev = bytes(len(rows[0]))
    for y, raw in enumerate(rows):
        f = filters[y % len(filters)]
        line = bytearray()
        for i, v in enumerate(raw):
            a = raw[i - bpp] if i >= bpp else 0
            b = prev[i]
            c = prev[i - bpp] if i >= bpp else 0
            if f == 0:
                p = 0
            elif f == 1:
                p = a
            elif f == 2:
                p = b
            elif f == 3:
                p = (a + b) // 2
            else:
                q = a + b - c
                pa, pb, pc = abs(q - a), abs(q - b), abs(q - c)
                p = a if pa <= pb and pa <= pc else (b if pb <= pc else c)
            line.append((v - p) & 255)
        out += bytes([f]) + bytes(line)
        prev = raw

    def chunk(tag, data):
        return struct.pack(">I", len(data)) + tag + data + struct.pack(">I", zlib.crc32(tag + data) & 0xFFFFFFFF)

    z = zlib.compress(out, 6)
    body = chunk(b"IHDR", struct.pack(">IIBBBBB", w, h, depth, colour, 0, 0, 0))
    if plte is not None:
        body += chunk(b"PLTE", bytes(plte))
    if trns is not None:
        body += chunk(b"tRNS", bytes(trns))
    body += chunk(b"IDAT", z[: len(z) // 2]) + chunk(b"IDAT", z[len(z) // 2:])  # two IDAT chunks
    return b"\x89PNG\r\n\x1a\n" + body + chunk(b"IEND", b"")


def _textured_tree(d):
    """seven materials, each with a PNG of another kind as map_Kd (and one bump map), on seven triangles"""
    rng = np.random.default_rng(11)
    os.makedirs(os.path.join(d, "objs"))
    kinds = [  # name, colour type, depth, channels, extra chunks
        ("rgba8", 6, 8, 4, {}), ("rgb8", 2, 8, 3, {}), ("grey8", 0, 8, 1, {}), ("ga8", 4, 8, 2, {}), ("rgb16", 2, 16, 3, {}),
        ("pal4", 3, 4, 1, {"plte": rng.integers(0, 256, 48).tolist(), "trns": rng.integers(0, 256, 9).tolist()}),
        ("grey2", 0, 2, 1, {"trns": [0, 2]}), ("rgbkey", 2, 8, 3, {"trns": [0, 7, 0, 7, 0, 7]}),
    ]
    mtl, obj = "", "mtllib m.mtl\nv 0 0 0\nv 1 0 0\nv 0 1 0\nvn 0 0 1\nvt 0 0\nvt 1 0\nvt 0 1\n"
    for i, (name, colour, depth, ch, extra) in enumerate(kinds):
        w, h = 5 + i, 7 - (i % 3)
        s = rng.integers(0, 1 << depth, (h, w, ch))
        if name == "rgbkey":
            s[1, 2] = 7  # a pixel equal to the colour key
        with open(os.path.join(d, "objs", name + ".png"), "wb") as f:
            f.write(_png_bytes(s, colour, depth, [0, 1, 2, 3, 4], **extra))
        mtl += f"newmtl m{i}\nKd 0.5 0.5 0.5\nmap_Kd {name}.png\n"
        if i == 1:
            mtl += "map_Bump -bm 0.3 rgba8.png\n"
        obj += f"usemtl m{i}\nf 1/1/1 2/2/1 3/3/1\n"
    with open(os.path.join(d, "objs", "m.mtl"), "w") as f:
        f.write(mtl)
    with open(os.path.join(d, "objs", "a.obj"), "w") as f:
        f.write(obj)
    with open(os.path.join(d, "s.scene"), "w") as f:
        f.write("[Models]\nfile: objs/a.obj\npos: 0,0,0\nscale: 1,1,1\n\n[Materials]\nname: m2\ncolor: 1,0,0\n\n[Camera]\npos: 0,0,-5\nforward: 0,0,1\nfocus: 1\n")
    return os.path.join(d, "s.scene")


def test_png_textures_decode(ptb, tmp_path):
    """8-bit RGBA texels survive the trip exactly; the override of m2 drops its texture like `mat.pbr = m.mat.pbr` does"""
    scene = _textured_tree(str(tmp_path))
    _convert(ptb, scene, str(tmp_path / "s.ptbs"), str(tmp_path))
    s = ptb.scenes.read_ptbs(str(tmp_path / "s.ptbs"))
    assert len(s["textures"]) == 8  # 8 albedo maps - the overridden one + the bump map
    bind = [(m.get("tex", -1), m.get("nrm", -1)) for m in s["materials"]]
    assert bind == [(-1, -1), (0, -1), (1, 2), (-1, -1), (3, -1), (4, -1), (5, -1), (6, -1), (7, -1)]
    assert (s["textures"][0] == s["textures"][2]).all() and s["textures"][0].shape == (7, 5, 4)
    assert float(s["textures"][1][..., 3].min()) == 1.0  # RGB: opaque
    assert (s["textures"][7][1, 2] == np.array([7 / 255, 7 / 255, 7 / 255, 0], np.float32)).all()  # the colour key


@pytest.mark.skipif(not have_ref(), reason="oracle/_ref/ptref not built (needs the reference sources)")
def test_png_textures_equal_stb_image_in_the_compiled_reference(ptb, tmp_path):
    scene = _textured_tree(str(tmp_path))
    mine = _convert(ptb, scene, str(tmp_path / "mine.ptbs"), str(tmp_path))
    run_ptref(["--scene-file", scene, "--base", str(tmp_path), "dump-scene", str(tmp_path / "ref.ptbs")])
    with open(tmp_path / "ref.ptbs", "rb") as f:
        assert mine == f.read()


# ---- polygons of five and more corners: tinyobj's built-in ear clipping ------------------------------------------------------
def _polygon_tree(d):
    rng = np.random.default_rng(5)
    lines, nv = ["mtllib m.mtl", "vn 0 0 1", "vt 0.5 0.5", "usemtl a"], 0

    def poly(pts):
        nonlocal nv
        lines.extend("v %.6f %.6f %.6f" % tuple(p) for p in pts)
        lines.append("f " + " ".join("%d/1/1" % (nv + i + 1) for i in range(len(pts))))
        nv += len(pts)

    poly([(np.cos(a), np.sin(a), 0.0) for a in np.linspace(0, 2 * np.pi, 6)[:-1]])          # convex pentagon
    poly([(0, 0, 1), (2, 0, 1), (2, 1, 1), (1, 1, 1), (1, 2, 1), (0, 2, 1)])                # concave L
    poly([(3, 0, 0), (3, 0, 2), (3, 1, 1), (3, 2, 2), (3, 2, 0), (3, 1, 0.5), (3, 0.5, 0.2)])  # clockwise, in the yz plane
    ang, r = np.linspace(0, 2 * np.pi, 11)[:-1], np.where(np.arange(10) % 2 == 0, 2.0, 0.8)
    poly([(r[i] * np.cos(ang[i]), r[i] * np.sin(ang[i]), 0.3 * r[i] * np.cos(ang[i]) + 5) for i in range(10)])  # tilted star
    for t in range(40):  # star-shaped polygons of 5..12 corners in random planes, both windings
        k = int(rng.integers(5, 13))
        a, rad = np.sort(rng.uniform(0, 2 * np.pi, k)), rng.uniform(0.3, 1.5, k)
        u = rng.normal(size=3)
        u /= np.linalg.norm(u)
        w = np.cross(u, rng.normal(size=3))
        w /= np.linalg.norm(w)
        c = rng.uniform(-5, 5, 3)
        pts = [c + rad[i] * (np.cos(a[i]) * u + np.sin(a[i]) * w) for i in range(k)]
        poly(pts[::-1] if t % 2 else pts)
    poly([(0, 0, 9), (1, 0, 9), (2, 0, 9), (3, 0, 9), (4, 0, 9)])  # collinear: no ear is ever found
    os.makedirs(os.path.join(d, "objs"))
    with open(os.path.join(d, "objs", "a.obj"), "w") as f:
        f.write("\n".join(lines) + "\n")
    with open(os.path.join(d, "objs", "m.mtl"), "w") as f:
        f.write("newmtl a\nKd 1 0 0\n")
    with open(os.path.join(d, "s.scene"), "w") as f:
        f.write("[Models]\nfile: objs/a.obj\npos: 0,0,0\nscale: 1,1,1\n")
    return os.path.join(d, "s.scene")


def test_polygons_are_ear_clipped(ptb, tmp_path):
    scene = _polygon_tree(str(tmp_path))
    _convert(ptb, scene, str(tmp_path / "s.ptbs"), str(tmp_path))
    p = ptb.scenes.read_ptbs(str(tmp_path / "s.ptbs"))["pos"].astype(np.float64).reshape(-1, 3, 3)
    area = 0.5 * np.linalg.norm(np.cross(p[:, 1] - p[:, 0], p[:, 2] - p[:, 0]), axis=1)
    # pentagon -> 3 triangles covering 5/2 sin(72 deg); the L -> 4 triangles of total area 3
    assert np.isclose(area[:3].sum(), 2.5 * np.sin(2 * np.pi / 5), rtol=1e-5) and np.isclose(area[3:7].sum(), 3.0, rtol=1e-6)
    assert (p[3:7, :, 2] == 1).all() and (p[7:12, :, 0] == 3).all()  # 7 corners -> 5 triangles in x = 3


@pytest.mark.skipif(not have_ref(), reason="oracle/_ref/ptref not built (needs the reference sources)")
def test_polygons_equal_the_compiled_reference(ptb, tmp_path):
    scene = _polygon_tree(str(tmp_path))
    mine = _convert(ptb, scene, str(tmp_path / "mine.ptbs"), str(tmp_path))
    run_ptref(["--scene-file", scene, "--base", str(tmp_path), "dump-scene", str(tmp_path / "ref.ptbs")])
    with open(tmp_path / "ref.ptbs", "rb") as f:
        assert mine == f.read()


# ---- TGA textures (decode_tga) against stb_image in the compiled reference ---------------------------------------------------
def _tga_bytes(px, rle, top_down):
    """px: (H, W, C) uint8 in RGB(A) / grey order; the file stores BGR(A), rows bottom-up unless top_down"""
    import struct
    h, w, c = px.shape
    rows = px if top_down else px[::-1]
    data = rows[..., [2, 1, 0, 3][:c]] if c >= 3 else rows
    flat = np.ascontiguousarray(data, np.uint8).reshape(-1, c)
    hdr = struct.pack("<BBBHHBHHHHBB", 0, 0, (3 if c == 1 else 2) + (8 if rle else 0), 0, 0, 0, 0, 0, w, h, 8 * c, (0x20 if top_down else 0) | (8 if c == 4 else 0))
    if not rle:
        return hdr + flat.tobytes()
    out, i, n = b"", 0, len(flat)
    while i < n:  # alternate run packets and raw packets, crossing row ends on purpose
        run = 1
        while i + run < n and run < 128 and (flat[i + run] == flat[i]).all():
            run += 1
        if run > 1:
            out += bytes([0x80 | (run - 1)]) + flat[i].tobytes()
            i += run
        else:
            k = min(5, n - i)
            out += bytes([k - 1]) + flat[i:i + k].tobytes()
            i += k
    return hdr + out


def _tga_tree(d):
    rng = np.random.default_rng(21)
    os.makedirs(os.path.join(d, "objs"))
    mtl, obj = "", "mtllib m.mtl\nv 0 0 0\nv 1 0 0\nv 0 1 0\nvn 0 0 1\nvt 0 0\nvt 1 0\nvt 0 1\n"
    for i, (c, rle, top) in enumerate([(3, False, False), (4, False, True), (1, False, False), (3, True, False), (4, True, True), (1, True, True)]):
        px = rng.integers(0, 4, (6 + i, 9 - i, c)).astype(np.uint8) * 60  # few values: real runs for the RLE flavours
        with open(os.path.join(d, "objs", f"t{i}.tga"), "wb") as f:
            f.write(_tga_bytes(px, rle, top))
        mtl += f"newmtl m{i}\nKd 0.5 0.5 0.5\nmap_Kd t{i}.tga\n"
        obj += f"usemtl m{i}\nf 1/1/1 2/2/1 3/3/1\n"
    with open(os.path.join(d, "objs", "m.mtl"), "w") as f:
        f.write(mtl)
    with open(os.path.join(d, "objs", "a.obj"), "w") as f:
        f.write(obj)
    with open(os.path.join(d, "s.scene"), "w") as f:
        f.write("[Models]\nfile: objs/a.obj\npos: 0,0,0\nscale: 1,1,1\n")
    return os.path.join(d, "s.scene")


def test_tga_textures_decode(ptb, tmp_path):
    scene = _tga_tree(str(tmp_path))
    _convert(ptb, scene, str(tmp_path / "s.ptbs"), str(tmp_path))
    tex = ptb.scenes.read_ptbs(str(tmp_path / "s.ptbs"))["textures"]
    rng = np.random.default_rng(21)
    for i, c in enumerate([3, 4, 1, 3, 4, 1]):  # the generator's pixels, row 0 = top, whatever the file's row order / packing
        px = rng.integers(0, 4, (6 + i, 9 - i, c)).astype(np.uint8) * 60
        want = np.full((6 + i, 9 - i, 4), 255, np.uint8)
        want[..., :3] = px[..., :3] if c >= 3 else px
        if c == 4:
            want[..., 3] = px[..., 3]
        assert (tex[i] == want.astype(np.float32) / np.float32(255)).all(), i


@pytest.mark.skipif(not have_ref(), reason="oracle/_ref/ptref not built (needs the reference sources)")
def test_tga_textures_equal_stb_image_in_the_compiled_reference(ptb, tmp_path):
    scene = _tga_tree(str(tmp_path))
    mine = _convert(ptb, scene, str(tmp_path / "mine.ptbs"), str(tmp_path))
    run_ptref(["--scene-file", scene, "--base", str(tmp_path), "dump-scene", str(tmp_path / "ref.ptbs")])
    with open(tmp_path / "ref.ptbs", "rb") as f:
        assert mine == f.read()


def test_group_of_only_degenerate_faces_is_skipped(ptb, tmp_path):
    """A group whose faces are all dropped (fewer than three corners) still becomes a shape in tinyobjloader, with an empty
    mesh; the reference reads material_ids[0] of that empty vector (object.cpp:92).  The loader emits no object for it."""
    d = str(tmp_path)
    os.makedirs(os.path.join(d, "objs"))
    with open(os.path.join(d, "objs", "a.obj"), "w") as f:
        f.write("v 0 0 0\nv 1 0 0\nv 1 1 0\nvn 0 0 1\nvt 0 0\ng broken\nf 1/1/1 2/1/1\nf 3/1/1\ng good\nf 1/1/1 2/1/1 3/1/1\n")
    with open(os.path.join(d, "s.scene"), "w") as f:
        f.write("[Models]\nfile: objs/a.obj\n")
    ptb.convert_scene_file(os.path.join(d, "s.scene"), os.path.join(d, "o.ptbs"), base_dir=d)
    s = ptb.scenes.read_ptbs(os.path.join(d, "o.ptbs"))
    assert s["pos"].shape == (1, 9)
    assert len(s["objects"]) == 1 and s["objects"][0]["end"] - s["objects"][0]["start"] == 1
