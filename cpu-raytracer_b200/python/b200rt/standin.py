"""Deterministic procedural stand-ins for the assets that are missing from the reference checkout.

`Data/Sponza/sponza.obj` (262 267 triangles) and `Data/sibenik/sibenik.obj` (75 284) are listed in
the reference's .MISSING_LARGE_BLOBS and cannot be fetched (no network), so the Sponza / Sibenik
benchmark configurations run on generated geometry of matching triangle count and similar
character: a long colonnaded hall, coarse floor and walls, finely tessellated columns, arches and
hanging cloth (non-uniform triangle density, long thin triangles, large occluders), with 1024^2
textures. Also generated: a biconvex glass lens and a curved mirror (the roles Magnifier.obj and
Concave.obj play in SCENE_SPONZA, Scene.cpp:75-130) and the sky probe. Everything is written as
ordinary OBJ / MTL / TGA / .float files, so the host library AND the reference binary load them
through their normal loaders. Every result on these assets must be labelled "stand-in".
"""
import os

import numpy as np

from . import scenes


class MeshBuilder:
    """Accumulates indexed triangles with per-vertex normal / uv and a material per face."""

    def __init__(self):
        self.v, self.vn, self.vt, self.faces, self.face_mat = [], [], [], [], []
        self.materials = []  # (name, lines)
        self._count = 0

    def material(self, name, kd=(1, 1, 1), ks=(0, 0, 0), kt=None, ni=1.0, map_kd=None):
        lines = ["newmtl " + name, "Kd %.6f %.6f %.6f" % tuple(kd), "Ks %.6f %.6f %.6f" % tuple(ks)]
        if kt is not None:
            lines.append("Kt %.6f %.6f %.6f" % tuple(kt))
        lines.append("Ni %.6f" % ni)
        if map_kd:
            lines.append("map_Kd " + map_kd)
        self.materials.append((name, lines))
        return len(self.materials) - 1

    def grid(self, position_fn, nu, nv, mat, uv_scale=(1.0, 1.0), flip=False):
        """Tessellates a parametric patch p(u, v), u, v in [0, 1], into nu x nv quads. Normals by finite differences."""
        u, v = np.meshgrid(np.linspace(0, 1, nu + 1), np.linspace(0, 1, nv + 1), indexing="ij")
        p = position_fn(u, v)
        e = 1e-3
        du = position_fn(np.clip(u + e, 0, 1), v) - position_fn(np.clip(u - e, 0, 1), v)
        dv = position_fn(u, np.clip(v + e, 0, 1)) - position_fn(u, np.clip(v - e, 0, 1))
        n = np.cross(du, dv)
        n /= np.maximum(np.linalg.norm(n, axis=-1, keepdims=True), 1e-20)
        if flip:
            n = -n
        uv = np.stack([u * uv_scale[0], v * uv_scale[1]], -1)

        base = self._count
        self.v.append(p.reshape(-1, 3)); self.vn.append(n.reshape(-1, 3)); self.vt.append(uv.reshape(-1, 2))
        self._count += (nu + 1) * (nv + 1)

        i, j = np.meshgrid(np.arange(nu), np.arange(nv), indexing="ij")
        a = base + i * (nv + 1) + j
        b = a + (nv + 1)
        c = b + 1
        d = a + 1
        t1 = np.stack([a, b, c], -1).reshape(-1, 3)
        t2 = np.stack([a, c, d], -1).reshape(-1, 3)
        tris = np.concatenate([t1, t2])
        if flip:
            tris = tris[:, ::-1]
        self.faces.append(tris)
        self.face_mat.append(np.full(len(tris), mat, dtype=np.int32))

    @property
    def triangle_count(self):
        return int(sum(len(f) for f in self.faces))

    def write(self, obj_path):
        v, vn, vt = np.concatenate(self.v), np.concatenate(self.vn), np.concatenate(self.vt)
        faces, mats = np.concatenate(self.faces) + 1, np.concatenate(self.face_mat)
        mtl_name = os.path.splitext(os.path.basename(obj_path))[0] + ".mtl"
        with open(os.path.splitext(obj_path)[0] + ".mtl", "w") as f:
            for _, lines in self.materials:
                f.write("\n".join(lines) + "\n\n")
        with open(obj_path, "w") as f:
            f.write("mtllib %s\n" % mtl_name)
            f.write("".join("v %.6f %.6f %.6f\n" % tuple(r) for r in v))
            f.write("".join("vt %.6f %.6f\n" % tuple(r) for r in vt))
            f.write("".join("vn %.6f %.6f %.6f\n" % tuple(r) for r in vn))
            order = np.argsort(mats, kind="stable")
            current = -1
            chunks = []
            for k in order:
                if mats[k] != current:
                    current = mats[k]
                    chunks.append("usemtl %s\n" % self.materials[current][0])
                a, b, c = faces[k]
                chunks.append("f %d/%d/%d %d/%d/%d %d/%d/%d\n" % (a, a, a, b, b, b, c, c, c))
            f.write("".join(chunks))
        return obj_path


# ------------------------------------------------------------------------------------ textures ---
def _write_tga(path, rgb):
    h, w, _ = rgb.shape
    header = bytearray(18)
    header[2] = 2
    header[12], header[13] = w & 255, w >> 8
    header[14], header[15] = h & 255, h >> 8
    header[16], header[17] = 24, 0x20
    with open(path, "wb") as f:
        f.write(bytes(header))
        f.write(np.ascontiguousarray(rgb[..., ::-1]).astype(np.uint8).tobytes())


def _noise(rng, size, octaves=5):
    img = np.zeros((size, size))
    for o in range(octaves):
        n = 4 << o
        coarse = rng.random((n, n))
        img += np.kron(coarse, np.ones((size // n, size // n))) / (1 << o)
    return img / img.max()


def make_textures(directory, seed, size=1024):
    rng = np.random.default_rng(seed)
    os.makedirs(directory, exist_ok=True)
    y, x = np.mgrid[0:size, 0:size]
    out = {}

    def save(name, rgb):
        path = os.path.join(directory, name + ".tga")
        _write_tga(path, np.clip(rgb, 0, 255))
        out[name] = path

    n = _noise(rng, size)
    brick = ((x // 128 + (y // 64) % 2 * 0.5) % 1 > 0.06) & ((y % 64) > 4)
    save("bricks", np.where(brick[..., None], np.stack([150 + 60 * n, 80 + 40 * n, 60 + 30 * n], -1), 190.0))
    tile = ((x % 256) > 6) & ((y % 256) > 6)
    check = ((x // 256 + y // 256) % 2)[..., None]
    save("floor", np.where(tile[..., None], (150 + 60 * check) + 40 * n[..., None] * np.ones(3), 60.0))
    save("stone", np.stack([170 + 60 * n, 165 + 55 * n, 150 + 50 * n], -1))
    save("ceiling", np.stack([120 + 50 * n, 110 + 40 * n, 100 + 40 * n], -1))
    weave = 0.75 + 0.25 * np.sin(x * 0.8) * np.sin(y * 0.8)
    for name, colour in (("fabric_red", (200, 40, 40)), ("fabric_green", (40, 160, 60)), ("fabric_blue", (50, 70, 200))):
        save(name, weave[..., None] * np.array(colour) * (0.7 + 0.3 * n[..., None]))
    return out


# ------------------------------------------------------------------------------------ geometry ---
def _cylinder(cx, cz, y0, y1, radius, flute=0.0, flutes=12):
    def fn(u, v):
        ang = u * 2 * np.pi
        r = radius * (1.0 + flute * np.cos(flutes * ang)) * (1.0 + 0.08 * np.cos(np.pi * (v - 0.5)) ** 8)
        return np.stack([cx + r * np.cos(ang), y0 + (y1 - y0) * v, cz + r * np.sin(ang)], -1)
    return fn


def hall(obj_path, texture_dir, seed, target_triangles, length=260.0, width=110.0, height=100.0):
    """A colonnaded hall centred on the origin, long axis = x, floor at y = 0."""
    tex = make_textures(texture_dir, seed)
    rel = lambda p: os.path.relpath(p, os.path.dirname(obj_path))
    m = MeshBuilder()
    bricks = m.material("bricks", map_kd=rel(tex["bricks"]))
    floor = m.material("floor", map_kd=rel(tex["floor"]))
    stone = m.material("stone", map_kd=rel(tex["stone"]))
    ceiling = m.material("ceiling", map_kd=rel(tex["ceiling"]))
    fabrics = [m.material(n, map_kd=rel(tex[n])) for n in ("fabric_red", "fabric_green", "fabric_blue")]

    hx, hz = length / 2, width / 2
    aisle = hz * 0.55  # columns stand at z = +-aisle

    # coarse shell: floor, walls, gallery floors, roof ring (open to the sky above the nave)
    m.grid(lambda u, v: np.stack([-hx + length * u, 0 * u, -hz + width * v], -1), 40, 20, floor, (20, 8), flip=True)
    for sign in (-1, 1):
        m.grid(lambda u, v, s=sign: np.stack([-hx + length * u, height * v, s * hz + 0 * u], -1), 30, 12, bricks, (16, 6), flip=(sign > 0))
        m.grid(lambda u, v, s=sign: np.stack([s * hx + 0 * u, height * v, -hz + width * u], -1), 14, 12, bricks, (8, 6), flip=(sign < 0))
        for level in (36.0, 68.0):  # gallery floors between wall and colonnade
            m.grid(lambda u, v, s=sign, y=level: np.stack([-hx + length * u, y + 0 * u, s * (aisle + (hz - aisle) * v)], -1), 30, 4, ceiling, (16, 2), flip=(sign < 0))
        m.grid(lambda u, v, s=sign: np.stack([-hx + length * u, height + 0 * u, s * (aisle * 0.8 + (hz - aisle * 0.8) * v)], -1), 30, 4, ceiling, (16, 2), flip=(sign < 0))

    coarse = m.triangle_count
    # fine detail: columns (three storeys), arches between them, cloth; resolution chosen to hit the target
    n_columns = 13
    xs = np.linspace(-hx * 0.88, hx * 0.88, n_columns)
    storeys = [(0.0, 36.0, 3.2), (36.0, 68.0, 2.4), (68.0, 100.0, 1.8)]
    n_cloth = 8
    detail_budget = target_triangles - coarse
    # budget split: columns 62 %, arches 14 %, cloth 24 %
    per_column = 0.62 * detail_budget / (2 * n_columns * len(storeys))
    seg = max(8, int(round(np.sqrt(per_column / 2 / 0.75))))     # seg around, 0.75*seg rings
    rings = max(6, int(round(0.75 * seg)))
    per_arch = 0.14 * detail_budget / (2 * (n_columns - 1) * len(storeys))
    arch_u = max(6, int(round(per_arch / 2 / 4)))
    per_cloth = 0.24 * detail_budget / n_cloth
    cloth_n = max(8, int(round(np.sqrt(per_cloth / 2))))

    for sign in (-1, 1):
        for (y0, y1, radius) in storeys:
            for cx in xs:
                m.grid(_cylinder(cx, sign * aisle, y0, y1 - 4.0, radius, flute=0.04), seg, rings, stone, (4, 6))
            for a, b in zip(xs[:-1], xs[1:]):
                def arch(u, v, a=a, b=b, y1=y1, s=sign):
                    ang = np.pi * u
                    cxm, r = 0.5 * (a + b), 0.5 * (b - a)
                    return np.stack([cxm - r * np.cos(ang), (y1 - 4.0 - r * 0.45) + r * 0.45 * np.sin(ang) + 0 * v, s * aisle + (v - 0.5) * 3.0], -1)
                m.grid(arch, arch_u, 4, stone, (6, 1))

    rng = np.random.default_rng(seed + 1)
    for k in range(n_cloth):
        cx = -hx * 0.75 + (length * 0.75) * k / (n_cloth - 1)
        phase, amp = rng.random() * 6.28, 1.5 + rng.random() * 1.5
        side = -1 if k % 2 else 1
        def cloth(u, v, cx=cx, phase=phase, amp=amp, side=side):
            x = cx + (u - 0.5) * 22.0
            y = 66.0 - 30.0 * v
            z = side * (aisle - 6.0) + amp * np.sin(u * 18.0 + phase) * (0.3 + v) + 0.8 * np.sin(v * 9.0 + phase)
            return np.stack([x, y, z], -1)
        m.grid(cloth, cloth_n, cloth_n, fabrics[k % 3], (3, 3))

    m.write(obj_path)
    return m.triangle_count


def lens(obj_path, radius=5.5, thickness=1.4, n=48):
    """Biconvex glass lens in the y-z plane (optical axis = x): Kt 1, Ni 1.45, Ks 0.5 like Magnifier.mtl's Lense."""
    m = MeshBuilder()
    glass = m.material("Lense", kd=(0, 0, 0), ks=(0.5, 0.5, 0.5), kt=(1, 1, 1), ni=1.45)
    for sign in (-1, 1):
        def cap(u, v, s=sign):
            r = radius * v
            ang = 2 * np.pi * u
            bulge = s * 0.5 * thickness * (1.0 - v * v)
            return np.stack([bulge, r * np.cos(ang), r * np.sin(ang)], -1)
        m.grid(cap, n, n // 2, glass, flip=(sign > 0))
    m.write(obj_path)
    return m.triangle_count


def mirror(obj_path, size=10.0, depth=1.5, n=24):
    """Slightly concave mirror facing +x (Ks 1, like Concave.mtl)."""
    m = MeshBuilder()
    metal = m.material("Mirror", kd=(0, 0, 0), ks=(1, 1, 1))
    def sheet(u, v):
        y, z = (v - 0.0) * size, (u - 0.5) * size
        return np.stack([-depth * (1.0 - ((u - 0.5) * 2) ** 2) * (1.0 - ((v - 0.5) * 2) ** 2), y, z], -1)
    m.grid(sheet, n, n, metal)
    m.write(obj_path)
    return m.triangle_count


# -------------------------------------------------------------------------------------- scenes ---
def cache_dir():
    return os.environ.get("B200RT_CACHE", "/tmp/b200rt_cache")


def _ensure(directory, name, seed, triangles):
    os.makedirs(directory, exist_ok=True)
    stamp = os.path.join(directory, name + ".ok")
    hall_obj = os.path.join(directory, name + "/hall.obj")
    if not os.path.exists(stamp):
        os.makedirs(os.path.dirname(hall_obj), exist_ok=True)
        count = hall(hall_obj, os.path.join(directory, name, "textures"), seed, triangles)
        lens(os.path.join(directory, name, "lens.obj"))
        mirror(os.path.join(directory, name, "mirror.obj"))
        scenes.write_sky(os.path.join(directory, name, "sky_256.float"), 256)
        with open(stamp, "w") as f:
            f.write(str(count))
    return os.path.join(directory, name)


def sponza_standin(out=None, width=1920, height=1080, directory=None):
    """Config 2: SCENE_SPONZA's set-up (Scene.cpp:75-130) on the stand-in hall (seed 1234, ~262 k triangles),
    plus the point light the benchmark configuration adds."""
    d = _ensure(directory or cache_dir(), "sponza_standin_v1", 1234, 262267)
    s = scenes.SceneWriter(width, height)
    s.camera((19.729143, 18.946165, 0.0), (0.0, -0.707107, 0.0, 0.707107))
    s.ambient((0.2, 0.2, 0.2))
    s.sky(os.path.join(d, "sky_256.float"))
    s.mesh(os.path.join(d, "hall.obj"))
    s.mesh(os.path.join(d, "lens.obj"), (0.0, 18.9, 0.0))
    s.mesh(os.path.join(d, "mirror.obj"), (20.0, 0.0, 0.0), scenes.axis_angle((0, 1, 0), scenes.PI))
    s.point_light((2000.0, 1800.0, 1600.0), (0.0, 60.0, 0.0))
    s.directional_light((0.9, 0.9, 0.9), _normalize((0.1, -1.0, 0.1)))
    return s.write(out or os.path.join(d, "sponza_%dx%d.scene" % (width, height)))


def sibenik_standin(out=None, width=3840, height=2160, directory=None):
    """Config 5: cathedral-like nave stand-in (seed 4321, ~75 k triangles), spot + point light."""
    d = _ensure(directory or cache_dir(), "sibenik_standin_v1", 4321, 75284)
    s = scenes.SceneWriter(width, height)
    s.camera((-100.0, 22.0, 4.0), scenes.axis_angle((0, 1, 0), scenes.PI * scenes.F(0.5)))
    s.ambient((0.2, 0.2, 0.2))
    s.sky(os.path.join(d, "sky_256.float"))
    s.mesh(os.path.join(d, "hall.obj"))
    s.mesh(os.path.join(d, "mirror.obj"), (60.0, 5.0, 0.0))
    s.point_light((2500.0, 2200.0, 2000.0), (-20.0, 50.0, 0.0))
    s.spot_light((4000.0, 4000.0, 3500.0), (40.0, 80.0, 0.0), _normalize((0.2, -1.0, 0.1)), 70.0, 80.0)
    return s.write(out or os.path.join(d, "sibenik_%dx%d.scene" % (width, height)))


def dielectrics_hall(data, out=None, width=1920, height=1080, directory=None):
    """Config 4 as SURVEY.md 8(d) states it: the Sponza hall (stand-in, seed 1234) with the reference's own
    Magnifier.obj in front of the shipped camera (Lense: Kt 1, Ni 1.45, Ks 0.5 — at the origin, where SCENE_SPONZA puts
    it, Scene.cpp:75-81), Diamond.obj (Kt (0.5, 0.1, 0.5), Ni 2.4) and the two glass spheres of Scene.cpp:8-19 (ior 1.33 /
    1.68, transmittance 0.6) behind the lens, Concave.obj where SCENE_SPONZA has it, the shipped directional light and
    the point light of config 2. 8 bounces, EWA, MAX_ANISOTROPY 8 are the renderer's settings. `data` holds the
    reference's shipped meshes (oracle/_ref/Data)."""
    d = _ensure(directory or cache_dir(), "sponza_standin_v1", 1234, 262267)
    s = scenes.SceneWriter(width, height)
    s.camera((19.729143, 18.946165, 0.0), (0.0, -0.707107, 0.0, 0.707107))   # Scene.cpp:127-128
    s.ambient((0.2, 0.2, 0.2))
    s.sky(os.path.join(d, "sky_256.float"))
    s.sphere(2.0, (-7.0, 16.0, -4.5), diffuse=(0.2, 0.2, 0.0), reflection=(0.6, 0.6, 0.0), transmittance=(0.6, 0.6, 0.6), ior=1.33)
    s.sphere(2.0, (-7.0, 16.0, +4.5), diffuse=(0.0, 0.2, 0.2), reflection=(0.0, 0.6, 0.6), transmittance=(0.6, 0.6, 0.6), ior=1.68)
    s.mesh(os.path.join(d, "hall.obj"))
    s.mesh(os.path.join(data, "Magnifier.obj"))
    s.mesh(os.path.join(data, "Diamond.obj"), (-5.0, 20.0, 0.0), scenes.axis_angle((0, 0, 1), scenes.F(0.6)))
    s.mesh(os.path.join(data, "Concave.obj"), (20.0, 0.0, 0.0), scenes.axis_angle((0, 1, 0), scenes.PI))
    s.point_light((2000.0, 1800.0, 1600.0), (0.0, 60.0, 0.0))
    s.directional_light((0.9, 0.9, 0.9), _normalize((0.1, -1.0, 0.1)))
    return s.write(out or os.path.join(d, "dielectrics_hall_%dx%d.scene" % (width, height)))


def _normalize(v):
    v = np.asarray(v, dtype=np.float64)
    return tuple(v / np.linalg.norm(v))
