"""Scene descriptions for the benchmark configurations (BASELINE.json `configs`).

A scene description is a small text file read by BOTH the host library (Scene::load) and the
headless build of the reference renderer (oracle/ref_recipe/headless/SceneFromFile.cpp), so the
two renderers are guaranteed to see the same scene. One item per line:

    b200rt_scene 1
    resolution W H
    camera FOV_DEG  px py pz  qx qy qz qw
    ambient r g b
    sky PATH                                  raw float RGB square probe (Sky.cpp:8-26)
    sphere RADIUS px py pz  <material>
    plane  px py pz  qx qy qz qw  <material>
    mesh   PATH.obj  px py pz  qx qy qz qw
    point_light        cr cg cb  px py pz
    spot_light         cr cg cb  px py pz  dx dy dz  INNER_DEG OUTER_DEG
    directional_light  cr cg cb  dx dy dz
    <material> = diffuse(3) reflection(3) transmittance(3) ior texture|-

Paths are absolute or relative to the scene file. Floats are written with 9 significant digits so
that they round-trip to the same binary32 value in every reader.
"""
import os

import numpy as np

F = np.float32
PI = F(3.14159265359)  # Util.h:8


def _f(x):
    return "%.9g" % float(F(x))


def _fs(values):
    return " ".join(_f(v) for v in values)


def axis_angle(axis, angle):
    """Quaternion::axis_angle in binary32 (Quaternion.h:28-38)."""
    half = F(0.5) * F(angle)
    s = F(np.sin(half, dtype=F))
    return [F(axis[0]) * s, F(axis[1]) * s, F(axis[2]) * s, F(np.cos(half, dtype=F))]


def rotate(q, v):
    """Quaternion * Vector3 (Quaternion.h:128-134), float64 is fine: the result is written as text."""
    u = np.array(q[:3], dtype=np.float64)
    v = np.array(v, dtype=np.float64)
    w = float(q[3])
    return 2.0 * np.dot(u, v) * u + (w * w - np.dot(u, u)) * v + 2.0 * w * np.cross(u, v)


def deg_to_rad(a):
    return F(a) * PI * F(0.00555555555)


def write_sky(path, size=256):
    """Procedural stand-in for the missing light probes: rgb = (0.4+0.4x/size, 0.6, 0.5+0.5y/size)."""
    y, x = np.mgrid[0:size, 0:size].astype(np.float32)
    data = np.stack([0.4 + 0.4 * x / size, np.full_like(x, 0.6), 0.5 + 0.5 * y / size], axis=-1).astype("<f4")
    data.tofile(path)
    return path


class SceneWriter:
    def __init__(self, width, height, fov_deg=110.0):
        self.lines = ["b200rt_scene 1", f"resolution {width} {height}"]
        self.fov = fov_deg

    def camera(self, position, rotation=(0, 0, 0, 1)):
        self.lines.append(f"camera {_f(self.fov)} {_fs(position)} {_fs(rotation)}")

    def ambient(self, rgb):
        self.lines.append(f"ambient {_fs(rgb)}")

    def sky(self, path):
        self.lines.append(f"sky {path}")

    @staticmethod
    def _material(diffuse=(1, 1, 1), reflection=(0, 0, 0), transmittance=(0, 0, 0), ior=1.0, texture=None):
        return f"{_fs(diffuse)} {_fs(reflection)} {_fs(transmittance)} {_f(ior)} {texture or '-'}"

    def sphere(self, radius, position, **material):
        self.lines.append(f"sphere {_f(radius)} {_fs(position)} {self._material(**material)}")

    def plane(self, position, rotation=(0, 0, 0, 1), **material):
        self.lines.append(f"plane {_fs(position)} {_fs(rotation)} {self._material(**material)}")

    def mesh(self, path, position=(0, 0, 0), rotation=(0, 0, 0, 1)):
        self.lines.append(f"mesh {path} {_fs(position)} {_fs(rotation)}")

    def point_light(self, colour, position):
        self.lines.append(f"point_light {_fs(colour)} {_fs(position)}")

    def spot_light(self, colour, position, direction, inner_deg, outer_deg):
        self.lines.append(f"spot_light {_fs(colour)} {_fs(position)} {_fs(direction)} {_f(inner_deg)} {_f(outer_deg)}")

    def directional_light(self, colour, direction):
        self.lines.append(f"directional_light {_fs(colour)} {_fs(direction)}")

    def write(self, path):
        os.makedirs(os.path.dirname(os.path.abspath(path)), exist_ok=True)
        partial = "%s.%d.part" % (path, os.getpid())       # several ranks may write the same scene file: never expose half of it
        with open(partial, "w") as f:
            f.write("\n".join(self.lines) + "\n")
        os.replace(partial, path)
        return path


# ------------------------------------------------------------------------------------------------
# Scenes built from the reference's shipped assets (Data/*.obj, *.png). `data` is a directory that
# holds them (a writable copy: `<obj>.bvh` caches are written next to the OBJ, BottomLevelBVH.cpp:26,53).

def monkey_plane(data, out, width=1280, height=720):
    """Config 1: Monkey.obj + textured plane + one point light (1 bounce, anisotropic filter)."""
    s = SceneWriter(width, height)
    s.camera((0.0, 1.0, -4.0))
    s.ambient((0.2, 0.2, 0.2))
    s.sky(os.path.join(data, "sky_256.float"))
    s.plane((0.0, -1.0, 0.0), reflection=(0.1, 0.1, 0.1), texture=os.path.join(data, "floor.png"))  # Scene.cpp:21-24
    s.mesh(os.path.join(data, "Monkey.obj"))
    s.point_light((10.0, 10.0, 10.0), (0.0, 5.0, -3.0))
    return s.write(out)


def dynamic(data, out, width=1920, height=1080):
    """Config 3: the shipped SCENE_DYNAMIC (Scene.cpp:7-71). Animate with animation=1."""
    s = SceneWriter(width, height)
    s.camera((-4.694016, 6.446100, -0.572288), (0.268476, 0.423740, -0.133092, 0.854779))
    s.ambient((0.2, 0.2, 0.2))
    s.sky(os.path.join(data, "sky_256.float"))
    s.sphere(1.0, (-2.0, 0.0, 10.0), diffuse=(0.2, 0.2, 0.0), reflection=(0.6, 0.6, 0.0), transmittance=(0.6, 0.6, 0.6), ior=1.33)
    s.sphere(1.0, (+2.0, 0.0, 10.0), diffuse=(0.0, 0.2, 0.2), reflection=(0.0, 0.6, 0.6), transmittance=(0.6, 0.6, 0.6), ior=1.68)
    s.plane((0.0, -1.0, 0.0), axis_angle((0, 1, 0), F(0.25) * PI), reflection=(0.1, 0.1, 0.1), texture=os.path.join(data, "floor.png"))
    for name, pos in [("Diamond", (0, 1, 0)), ("Monkey", (4, 2, 0)), ("icosphere", (0, 3, 4)), ("Rock", (6, 4, 4)), ("Torus", (0, 5, 8)), ("Torus", (-4, 2, 6))]:
        s.mesh(os.path.join(data, name + ".obj"), pos)
    s.point_light((0.0, 5.0, 10.0), (0.0, 0.0, 6.0))
    direction = rotate(axis_angle((1, 0, 0), deg_to_rad(70.0)), (0.0, 0.0, 1.0))
    s.spot_light((1.0, 0.0, 0.0), (0.0, 0.0, 10.0), direction, 70.0, 80.0)
    s.directional_light((0.5, 0.5, 0.5), (0.0, -1.0, 0.0))
    return s.write(out)


def dielectrics_small(data, out, width=1920, height=1080):
    """Config 4 without the (missing) Sponza hall: the camera looks through the lens of Magnifier.obj
    (Kt 1, Ni 1.45, placed as in SCENE_SPONZA, Scene.cpp:75-130) at Diamond.obj (Ni 2.4), two glass
    spheres (Scene.cpp:8-19), a Monkey, the Concave.obj mirror and a textured floor. 8 bounces and
    the EWA filter exercise the full reflect / refract recursion and the most expensive filter."""
    s = SceneWriter(width, height)
    s.camera((14.0, 19.0, 0.0), (0.0, -0.707107, 0.0, 0.707107))  # looks along -x, like the shipped camera
    s.ambient((0.2, 0.2, 0.2))
    s.sky(os.path.join(data, "sky_256.float"))
    s.sphere(2.0, (-7.0, 16.0, -4.5), diffuse=(0.2, 0.2, 0.0), reflection=(0.6, 0.6, 0.0), transmittance=(0.6, 0.6, 0.6), ior=1.33)
    s.sphere(2.0, (-7.0, 16.0, +4.5), diffuse=(0.0, 0.2, 0.2), reflection=(0.0, 0.6, 0.6), transmittance=(0.6, 0.6, 0.6), ior=1.68)
    s.plane((0.0, 12.0, 0.0), axis_angle((0, 1, 0), F(0.25) * PI), reflection=(0.1, 0.1, 0.1), texture=os.path.join(data, "bricks.png"))
    s.mesh(os.path.join(data, "Magnifier.obj"))
    s.mesh(os.path.join(data, "Diamond.obj"), (-5.0, 20.0, 0.0), axis_angle((0, 0, 1), F(0.6)))
    s.mesh(os.path.join(data, "Monkey.obj"), (-9.0, 22.5, 3.0), axis_angle((0, 1, 0), F(0.5) * PI))
    s.mesh(os.path.join(data, "Concave.obj"), (-16.0, 12.0, 0.0), axis_angle((0, 1, 0), PI))
    s.point_light((400.0, 400.0, 400.0), (6.0, 30.0, 6.0))
    s.directional_light((0.5, 0.5, 0.5), (-0.3, -1.0, 0.2))
    return s.write(out)
