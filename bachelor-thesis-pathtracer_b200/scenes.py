"""Scene containers and the synthetic workloads of BASELINE.json (host side, numpy).

* read_ptbs / write_ptbs: the flat `.ptbs` container of include/ptbs_format.h — what the
  reference holds in memory after initScene (src/scene/scene.cpp:97-131).
* lucy_standin: BASELINE configs 3/4.  showcase/objects/lucy.obj is not shipped with the reference
  (.MISSING_LARGE_BLOBS), so the statue is synthesised: an octahedron sphere subdivided 8 times
  (8 * 4^8 = 524,288 faces), stretched to statue proportions and displaced by seeded fractal
  noise (seed 0x1ACE), standing in the Cornell room with lucy.scene's material overrides and
  camera (showcase/scenes/lucy.scene:1-30).  SURVEY.md §8d, config 3.
* triangle_soup / soup_rays: BASELINE config 5: N triangles with constant expected
  occupancy and 2^k incoherent rays.
Everything is deterministic (fixed seeds), so the CPU reference and the GPU see identical bytes.
"""
import struct

import numpy as np

PTBS_MAGIC = 0x53425450
MAX_OBJ_MATS = 28
_OBJ = struct.Struct("<iifI28I64s")
_MAT = struct.Struct("<10f64s")
_HDR = struct.Struct("<8I8f")


def _name(b):
    return b.split(b"\0")[0].decode()


def read_ptbs(path):
    b = open(path, "rb").read()
    h = _HDR.unpack_from(b, 0)
    if h[0] != PTBS_MAGIC or h[1] not in (1, 2):
        raise ValueError("not a .ptbs v1/v2 container")
    n, no, nm = h[2], h[3], h[4]
    off = _HDR.size
    pos = np.frombuffer(b, np.float32, n * 9, off).reshape(n, 9).copy(); off += n * 36
    nrm = np.frombuffer(b, np.float32, n * 9, off).reshape(n, 9).copy(); off += n * 36
    uv = np.frombuffer(b, np.float32, n * 6, off).reshape(n, 6).copy(); off += n * 24
    mat = np.frombuffer(b, np.int32, n, off).copy(); off += n * 4
    objects = []
    for _ in range(no):
        f = _OBJ.unpack_from(b, off); off += _OBJ.size
        objects.append({"start": f[0], "end": f[1], "area": f[2], "materials": list(f[4:4 + f[3]]), "name": _name(f[32])})
    materials = []
    for _ in range(nm):
        f = _MAT.unpack_from(b, off); off += _MAT.size
        materials.append({"albedo": list(f[0:3]), "ior1": f[3], "ior2": f[4], "emission": f[5], "roughness": f[6],
                          "weights": list(f[7:10]), "name": _name(f[10])})
    cam = {"pos": list(h[8:11]), "forward": list(h[11:14]), "focus": h[14], "dof": h[15]}
    scene = {"pos": pos, "nrm": nrm, "uv": uv, "mat": mat, "objects": objects, "materials": materials, "camera": cam}
    if h[1] >= 2 and h[5] > 0:  # version 2: per-material {albedo texture, normal map} + the textures (ptbs_format.h)
        b2 = np.frombuffer(b, np.int32, 2 * nm, off).reshape(nm, 2); off += 8 * nm
        for m, (ta, tn) in zip(materials, b2):
            m["tex"], m["nrm"] = int(ta), int(tn)
        scene["textures"] = []
        for _ in range(h[5]):
            w, hh = struct.unpack_from("<ii", b, off); off += 8
            scene["textures"].append(np.frombuffer(b, np.float32, w * hh * 4, off).reshape(hh, w, 4).copy()); off += w * hh * 16
    return scene


def pack_objects(objects):
    out = bytearray()
    for o in objects:
        m = list(o["materials"])[:MAX_OBJ_MATS]
        out += _OBJ.pack(int(o["start"]), int(o["end"]), float(o["area"]), len(m), *(m + [0] * (MAX_OBJ_MATS - len(m))),
                         o["name"].encode()[:63])
    return bytes(out)


def pack_materials(materials):
    out = bytearray()
    for m in materials:
        out += _MAT.pack(*[float(x) for x in m["albedo"]], float(m["ior1"]), float(m["ior2"]), float(m["emission"]),
                         float(m["roughness"]), *[float(x) for x in m["weights"]], m["name"].encode()[:63])
    return bytes(out)


def material_bindings(scene):
    """n_materials x {albedo texture, normal map} (-1 = none) of a scene dict."""
    return np.array([[int(m.get("tex", -1)), int(m.get("nrm", -1))] for m in scene["materials"]], np.int32).reshape(-1, 2)


def write_ptbs(path, scene):
    n = scene["pos"].shape[0]
    c = scene["camera"]
    tex = scene.get("textures") or []
    with open(path, "wb") as f:
        f.write(_HDR.pack(PTBS_MAGIC, 2 if tex else 1, n, len(scene["objects"]), len(scene["materials"]), len(tex), 0, 0,
                          *[float(x) for x in c["pos"]], *[float(x) for x in c["forward"]], float(c["focus"]), float(c["dof"])))
        f.write(np.ascontiguousarray(scene["pos"], np.float32).tobytes())
        f.write(np.ascontiguousarray(scene["nrm"], np.float32).tobytes())
        f.write(np.ascontiguousarray(scene["uv"], np.float32).tobytes())
        f.write(np.ascontiguousarray(scene["mat"], np.int32).tobytes())
        f.write(pack_objects(scene["objects"]))
        f.write(pack_materials(scene["materials"]))
        if tex:
            f.write(material_bindings(scene).tobytes())
            for t in tex:
                t = np.ascontiguousarray(t, np.float32)
                f.write(struct.pack("<ii", t.shape[1], t.shape[0]))
                f.write(t.tobytes())


def upload_scene(pt, scene):
    """The uploads + camera of a scene dict through the C ABI."""
    pt.upload_triangles(scene["pos"], scene["nrm"], scene["uv"], scene["mat"])
    pt.upload_materials(pack_materials(scene["materials"]), len(scene["materials"]))
    pt.upload_objects(pack_objects(scene["objects"]), len(scene["objects"]))
    if scene.get("textures"):
        pt.upload_textures(scene["textures"], material_bindings(scene))
    c = scene["camera"]
    pt.set_camera(c["pos"], c["forward"], c["focus"], c["dof"])


# Bit-reproducibility: the generators use only +, -, *, /, sqrt, floor and comparisons on float64 (all
# correctly rounded by IEEE 754, no libm transcendental, no BLAS reduction), so every machine produces
# the same bytes and the committed reference digests (tests/golden/) stay valid on the GPU box.
def _norm3(v):
    return np.sqrt(v[..., 0] * v[..., 0] + v[..., 1] * v[..., 1] + v[..., 2] * v[..., 2])


def _cross(a, b):
    return np.stack([a[..., 1] * b[..., 2] - a[..., 2] * b[..., 1], a[..., 2] * b[..., 0] - a[..., 0] * b[..., 2],
                     a[..., 0] * b[..., 1] - a[..., 1] * b[..., 0]], -1)


def _psin(x):
    """Parabolic stand-in for sin(x) (period 2*pi, peaks +-1) from exactly rounded operations only."""
    t = x * (1.0 / 6.283185307179586)
    t = t - np.floor(t)
    return np.where(t < 0.5, 16.0 * t * (0.5 - t), -16.0 * (t - 0.5) * (1.0 - t))


def _tri_area(pos):
    p = pos.reshape(-1, 3, 3).astype(np.float64)
    return 0.5 * _norm3(_cross(p[:, 1] - p[:, 0], p[:, 2] - p[:, 0]))


def _flat_normals(pos):
    p = pos.reshape(-1, 3, 3).astype(np.float64)
    n = _cross(p[:, 1] - p[:, 0], p[:, 2] - p[:, 0])
    l = _norm3(n)[..., None]
    n = np.where(l > 0, n / np.maximum(l, 1e-300), np.array([0.0, 1.0, 0.0]))
    return np.repeat(n[:, None, :], 3, axis=1).reshape(-1, 9).astype(np.float32)


def _octa_sphere(level):
    """Unit octahedron sphere, every face subdivided into (2^level)^2 triangles -> [8*4^level, 3, 3] float64.
    Lattice points are exact multiples of 1/N, so vertices shared by faces are bit-identical."""
    N = 1 << level
    i, j = np.meshgrid(np.arange(N + 1), np.arange(N + 1), indexing="ij")
    # upward triangles (i,j),(i+1,j),(i,j+1) for i+j<N ; downward (i+1,j),(i+1,j+1),(i,j+1) for i+j<N-1
    iu, ju = np.nonzero((i + j < N) & (i < N) & (j < N))
    idn, jdn = np.nonzero((i + j < N - 1) & (i < N) & (j < N))
    a = np.concatenate([np.stack([iu, ju], 1), np.stack([idn + 1, jdn], 1)])
    b = np.concatenate([np.stack([iu + 1, ju], 1), np.stack([idn + 1, jdn + 1], 1)])
    c = np.concatenate([np.stack([iu, ju + 1], 1), np.stack([idn, jdn + 1], 1)])
    tri_ij = np.stack([a, b, c], 1).astype(np.float64)  # [N^2, 3, 2]
    faces = []
    for sx in (1.0, -1.0):
        for sy in (1.0, -1.0):
            for sz in (1.0, -1.0):
                A, B, C = np.array([sx, 0, 0]), np.array([0, sy, 0]), np.array([0, 0, sz])
                u, v = tri_ij[..., 0] / N, tri_ij[..., 1] / N
                w = 1.0 - u - v
                p = u[..., None] * A + v[..., None] * B + w[..., None] * C
                if sx * sy * sz < 0:  # keep outward winding
                    p = p[:, ::-1, :]
                faces.append(p)
    p = np.concatenate(faces)
    return p / _norm3(p)[..., None]


def _fractal(p, seed, octaves=6):
    """Seeded sum-of-waves fractal on unit directions p[..., 3] (float64), range about [-1, 1]."""
    rng = np.random.Generator(np.random.MT19937(seed))
    out = np.zeros(p.shape[:-1])
    amp, freq, tot = 1.0, 2.0, 0.0
    for _ in range(octaves):
        for _ in range(3):
            d = rng.uniform(-1.0, 1.0, size=3)
            d = d / np.sqrt(d[0] * d[0] + d[1] * d[1] + d[2] * d[2] + 1e-9)
            out += amp * _psin(freq * (p[..., 0] * d[0] + p[..., 1] * d[1] + p[..., 2] * d[2]) + rng.uniform(0.0, 6.0))
        tot += 3 * amp
        amp *= 0.5
        freq *= 2.0
    return out / tot


def statue_mesh(level=8, seed=0x1ACE, radius=(105.0, 225.0, 105.0), base=(0.0, -250.0, 0.0)):
    """[n,9] float32 positions of the stand-in statue standing on the Cornell floor."""
    s = _octa_sphere(level)
    r = 1.0 + 0.22 * _fractal(s, seed)
    # a waist/shoulder profile along y so the silhouette is not a plain ellipsoid
    prof = 0.75 + 0.25 * _psin(3.0 * s[..., 1] + 2.2)
    p = s * r[..., None]
    p[..., 0] *= radius[0] * prof
    p[..., 2] *= radius[2] * prof
    p[..., 1] = p[..., 1] * radius[1] + radius[1] * 1.0
    p += np.array(base)
    p[..., 1] += (base[1] + 1.0) - p[..., 1].min()  # stand on the floor
    return p.reshape(-1, 9).astype(np.float32)


def lucy_standin(room, level=8, seed=0x1ACE):
    """Scene dict: the Cornell room of `room` (a read_ptbs dict of cornel.ptbs) with its sphere replaced by
    the synthetic statue, lucy.scene's material overrides and camera."""
    keep = [o for o in room["objects"] if o["name"] != "Sphere"]
    pos, nrm, uv, mat, objects = [], [], [], [], []
    cur = 0
    for o in keep:
        n = o["end"] - o["start"]
        pos.append(room["pos"][o["start"]:o["end"]]); nrm.append(room["nrm"][o["start"]:o["end"]])
        uv.append(room["uv"][o["start"]:o["end"]]); mat.append(room["mat"][o["start"]:o["end"]])
        objects.append(dict(o, start=cur, end=cur + n))
        cur += n
    materials = [dict(m) for m in room["materials"]]
    # lucy.scene [Materials]: a named entry replaces the whole pbr + weights (scene.cpp:103-110;
    # unspecified refractive indices default to 1.0/1.5, shader.h:20-28)
    for name, w in (("Material.001", [0, 1, 0]), ("Material.002", [0, 1, 0])):
        for m in materials:
            if m["name"] == name:
                m.update(albedo=[1.0, 1.0, 1.0], emission=0.0, roughness=0.0, weights=w, ior1=1.0, ior2=1.5)
    materials.append({"albedo": [1.0, 0.4, 0.9], "ior1": 1.0, "ior2": 1.5, "emission": 0.0, "roughness": 0.1,
                      "weights": [1, 0, 0], "name": "Material.008"})
    sp = statue_mesh(level, seed)
    n = sp.shape[0]
    pos.append(sp); nrm.append(_flat_normals(sp)); uv.append(np.zeros((n, 6), np.float32))
    mat.append(np.full(n, len(materials) - 1, np.int32))
    objects.append({"start": cur, "end": cur + n, "area": float(_tri_area(sp).sum()), "materials": [len(materials) - 1],
                    "name": "LucyStandIn"})
    return {"pos": np.concatenate(pos), "nrm": np.concatenate(nrm), "uv": np.concatenate(uv), "mat": np.concatenate(mat),
            "objects": objects, "materials": materials,
            "camera": {"pos": [-250.0, 0.0, -180.0], "forward": [0.9, 0.0, 0.34], "focus": 1.0, "dof": 0.0}}


def triangle_soup(n, seed=None):
    """Config 5: triangle i has its centre uniform in [-100,100]^3 and vertices = centre + uniform[-s,s]^3 with
    s = 100 * n^(-1/3) (constant expected occupancy).  One object, one lambert material, no light."""
    seed = 7 + int(round(np.log10(n))) if seed is None else seed
    rng = np.random.Generator(np.random.MT19937(seed))
    s = 100.0 * float(n) ** (-1.0 / 3.0)
    c = rng.uniform(-100.0, 100.0, size=(n, 1, 3)).astype(np.float32)
    v = (c + rng.uniform(-s, s, size=(n, 3, 3)).astype(np.float32)).astype(np.float32)
    pos = v.reshape(n, 9)
    mats = [{"albedo": [0.8, 0.8, 0.8], "ior1": 1.0, "ior2": 1.5, "emission": 0.0, "roughness": 0.5, "weights": [1, 0, 0],
             "name": "Soup"}]
    return {"pos": pos, "nrm": _flat_normals(pos), "uv": np.zeros((n, 6), np.float32), "mat": np.zeros(n, np.int32),
            "objects": [{"start": 0, "end": n, "area": float(_tri_area(pos).sum()), "materials": [0], "name": "Soup"}],
            "materials": mats, "camera": {"pos": [0.0, 0.0, -400.0], "forward": [0.0, 0.0, 1.0], "focus": 1.0, "dof": 0.0}}


def _texture_bytes(w, h, salt, alpha):
    """Deterministic RGBA bytes -> Texture::data exactly as loadTexture leaves it (byte / 255.0f, texture.cpp:41-44).
    alpha: "opaque", or "holes" = a pattern of fully transparent, half transparent and opaque texels."""
    y, x = np.mgrid[0:h, 0:w].astype(np.int64)
    px = np.empty((h, w, 4), np.int64)
    for c in range(3):
        px[..., c] = (x * 37 + y * 91 + c * 53 + salt * 17 + ((x // 2 + y // 2) % 2) * 128) % 256
    if alpha == "opaque":
        px[..., 3] = 255
    else:
        px[..., 3] = np.choose((x + 2 * y) % 4, [255, 0, 128, 255])
    return (px.astype(np.float32) / np.float32(255.0)).astype(np.float32)


def textured_cornell(room, light_texture=True):
    """The Cornell box with textures on it — no textured scene ships with the reference, so this exercises the texture
    branches of the shader (shader.cpp:52-66,251-258) on its own scene: an albedo texture with see-through and
    half-transparent texels on two walls, a normal map on the floor and on the glass sphere, an opaque texture on the
    emissive cube (sampled by NEE), and uv stretched to [-1, 2] so that every branch of getTextureAtUV's wrap rule runs."""
    s = {k: (v.copy() if isinstance(v, np.ndarray) else v) for k, v in room.items()}
    s["materials"] = [dict(m) for m in room["materials"]]
    s["uv"] = (room["uv"].astype(np.float32) * np.float32(3.0) - np.float32(1.0)).astype(np.float32)
    s["textures"] = [_texture_bytes(16, 16, 1, "holes"), _texture_bytes(8, 5, 2, "opaque"), _texture_bytes(4, 4, 3, "opaque")]
    for mi, (ta, tn) in {1: (0, -1), 6: (0, 1), 5: (-1, 1), 4: (-1, 1), 7: (0, -1)}.items():
        s["materials"][mi]["tex"], s["materials"][mi]["nrm"] = ta, tn
    if light_texture:
        s["materials"][3]["tex"] = 2
    return s


def scene_from_triangles(pos):
    """A one-object, one-material lambert scene over explicit triangles (n x 9 floats; n may be 0) — edge-case tests."""
    pos = np.ascontiguousarray(pos, np.float32).reshape(-1, 9)
    n = pos.shape[0]
    mats = [{"albedo": [0.8, 0.8, 0.8], "ior1": 1.0, "ior2": 1.5, "emission": 0.0, "roughness": 0.5, "weights": [1, 0, 0],
             "name": "Plain"}]
    nrm = _flat_normals(pos) if n else np.zeros((0, 9), np.float32)
    area = float(_tri_area(pos).sum()) if n else 0.0
    return {"pos": pos, "nrm": nrm, "uv": np.zeros((n, 6), np.float32), "mat": np.zeros(n, np.int32),
            "objects": [{"start": 0, "end": n, "area": area, "materials": [0], "name": "Plain"}],
            "materials": mats, "camera": {"pos": [0.0, 0.0, -5.0], "forward": [0.0, 0.0, 1.0], "focus": 1.0, "dof": 0.0}}


def soup_rays(n, seed=99):
    """Config 5 rays: origin uniform on the sphere r = 250, direction towards a uniform point of [-100,100]^3
    (incoherent), normalised in float32.  [n, 6] float32."""
    rng = np.random.Generator(np.random.MT19937(seed))
    o = np.zeros((0, 3))
    while o.shape[0] < n:  # uniform directions by rejection from the cube (exact operations only)
        c = rng.uniform(-1.0, 1.0, size=(2 * n + 64, 3))
        r2 = c[:, 0] * c[:, 0] + c[:, 1] * c[:, 1] + c[:, 2] * c[:, 2]
        o = np.concatenate([o, c[(r2 <= 1.0) & (r2 > 1e-4)]])
    o = o[:n]
    o = (250.0 * o / _norm3(o)[:, None]).astype(np.float32)
    tgt = rng.uniform(-100.0, 100.0, size=(n, 3)).astype(np.float32)
    d = (tgt - o).astype(np.float32)
    d = (d / np.sqrt(d[:, 0:1] * d[:, 0:1] + d[:, 1:2] * d[:, 1:2] + d[:, 2:3] * d[:, 2:3])).astype(np.float32)
    return np.concatenate([o, d], 1).astype(np.float32)
