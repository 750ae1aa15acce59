"""Scene API kept verbatim from the reference (src/scene/__init__.py:1-3)."""
from .scene import Scene, Light, Plane, Sphere
from .rotation import euler_rotation
from .camera import Camera

__all__ = ["Scene", "Light", "Plane", "Sphere", "euler_rotation", "Camera"]
