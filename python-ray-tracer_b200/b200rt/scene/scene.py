"""Scene description and packing (API of reference src/scene/scene.py:9-115).

The packed arrays are the device-side ABI of the render path (include/b200rt.h):
    spheres float32 (7,S): rows Cx,Cy,Cz,r,R,G,B      lights float32 (3,L)
    planes  float32 (9,P): rows ox,oy,oz,nx,ny,nz,R,G,B with n normalised in float64 then stored float32
`generate_scene()` returns (spheres, lights, planes) in that order (scene.py:96-97).
"""
from __future__ import annotations
from dataclasses import dataclass
from typing import List
import numpy as np
from .colors import RED, BLUE, MAGENTA, YELLOW, GREEN, GREY
from .common import Vector3D, Color


@dataclass
class Sphere:
    origin: Vector3D
    radius: float
    color: Color
    data_length: int = 7

    def to_array(self) -> np.ndarray:
        out = np.zeros(self.data_length, dtype=np.float32)
        out[0:3] = np.array(self.origin)
        out[3] = self.radius
        out[4:7] = np.array(self.color)
        return out


@dataclass
class Light:
    origin: Vector3D
    data_length: int = 3

    def to_array(self) -> np.ndarray:
        out = np.zeros(self.data_length, dtype=np.float32)
        out[0:3] = np.array(self.origin)
        return out


@dataclass
class Plane:
    origin: Vector3D
    normal: Vector3D
    color: Color
    data_length: int = 9

    def to_array(self) -> np.ndarray:
        out = np.zeros(self.data_length, dtype=np.float32)
        n = np.array(self.normal)
        out[0:3] = np.array(self.origin)
        out[3:6] = n / np.linalg.norm(n)
        out[6:9] = np.array(self.color)
        return out


def _pack(objects, rows: int) -> np.ndarray:
    out = np.zeros((rows, len(objects)), dtype=np.float32)
    for i, obj in enumerate(objects):
        out[:, i] = obj.to_array()
    return out


class Scene:
    def __init__(self, lights: List[Light], spheres: List[Sphere], planes: List[Plane]):
        self.lights = lights
        self.spheres = spheres
        self.planes = planes

    def get_spheres(self) -> np.ndarray:
        return _pack(self.spheres, Sphere.data_length)

    def get_planes(self) -> np.ndarray:
        return _pack(self.planes, Plane.data_length)

    def get_lights(self) -> np.ndarray:
        return _pack(self.lights, Light.data_length)

    def generate_scene(self):
        return self.get_spheres(), self.get_lights(), self.get_planes()

    @classmethod
    def from_arrays(cls, sphere_centers=None, sphere_radii=None, sphere_colors=None, light_positions=None,
                    plane_points=None, plane_normals=None, plane_colors=None) -> "PackedScene":
        """Vectorised packing: (N,3) / (N,) arrays in, the same float32 (7,S)/(3,L)/(9,P) bytes out as
        building N dataclass objects and calling get_spheres()/get_lights()/get_planes() (scene.py:69-94), without the
        per-object Python loop."""
        return PackedScene(sphere_centers, sphere_radii, sphere_colors, light_positions, plane_points, plane_normals, plane_colors)

    @staticmethod
    def default_scene() -> "Scene":
        """The demo scene of reference scene.py:99-115: 3 lights, 6 spheres, 1 grey floor plane."""
        lights = [Light([2.5, -2.0, 3.0]), Light([2.5, 2.0, 3.0]), Light([5.0, 0.1, 6.0])]
        spheres = [Sphere([2.2, 0.3, 1.0], 1.0, RED), Sphere([0.6, 0.7, 0.4], 0.4, BLUE),
                   Sphere([0.6, -0.8, 0.5], 0.5, YELLOW), Sphere([-1.2, 0.2, 0.5], 0.5, MAGENTA),
                   Sphere([-1.7, -0.5, 0.3], 0.3, GREEN), Sphere([-2.0, 1.31, 1.3], 1.3, RED)]
        planes = [Plane([5, 0, 0], [0, 0, 1], GREY)]
        return Scene(lights, spheres, planes)


class PackedScene:
    """A scene held as packed arrays: quacks like Scene for the render path (generate_scene / get_*)."""

    def __init__(self, sphere_centers=None, sphere_radii=None, sphere_colors=None, light_positions=None,
                 plane_points=None, plane_normals=None, plane_colors=None):
        def mat(a, cols):
            a = np.zeros((0, cols)) if a is None else np.asarray(a, dtype=np.float64)
            return a.reshape(-1, cols)
        c, col = mat(sphere_centers, 3), mat(sphere_colors, 3)
        r = np.zeros(0) if sphere_radii is None else np.asarray(sphere_radii, dtype=np.float64).ravel()
        if not (len(c) == len(r) == len(col)):
            raise ValueError("sphere_centers, sphere_radii and sphere_colors must have the same length")
        self._spheres = np.zeros((7, len(c)), dtype=np.float32)
        self._spheres[0:3] = c.T; self._spheres[3] = r; self._spheres[4:7] = col.T
        li = mat(light_positions, 3)
        self._lights = np.ascontiguousarray(li.T, dtype=np.float32)
        po, pn, pc = mat(plane_points, 3), mat(plane_normals, 3), mat(plane_colors, 3)
        if not (len(po) == len(pn) == len(pc)):
            raise ValueError("plane_points, plane_normals and plane_colors must have the same length")
        self._planes = np.zeros((9, len(po)), dtype=np.float32)
        if len(po):
            nrm = np.array([np.linalg.norm(v) for v in pn])        # same routine per vector as Plane.to_array (scene.py:50-51)
            self._planes[0:3] = po.T; self._planes[3:6] = (pn / nrm[:, None]).T; self._planes[6:9] = pc.T

    def get_spheres(self) -> np.ndarray:
        return self._spheres.copy()

    def get_planes(self) -> np.ndarray:
        return self._planes.copy()

    def get_lights(self) -> np.ndarray:
        return self._lights.copy()

    def generate_scene(self):
        return self.get_spheres(), self.get_lights(), self.get_planes()
