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

    @staticmethod
    def default_scene() -> "Scene":
        """The demo scene of reference scene.py:99-115: 3 lights, 6 spheres, 1 grey floor plane."""
        lights = [Light([2.5, -2.0, 3.0]), Light([2.5, 2.0, 3.0]), Light([5.0, 0.1, 6.0])]
        spheres = [Sphere([2.2, 0.3, 1.0], 1.0, RED), Sphere([0.6, 0.7, 0.4], 0.4, BLUE),
                   Sphere([0.6, -0.8, 0.5], 0.5, YELLOW), Sphere([-1.2, 0.2, 0.5], 0.5, MAGENTA),
                   Sphere([-1.7, -0.5, 0.3], 0.3, GREEN), Sphere([-2.0, 1.31, 1.3], 1.3, RED)]
        planes = [Plane([5, 0, 0], [0, 0, 1], GREY)]
        return Scene(lights, spheres, planes)
