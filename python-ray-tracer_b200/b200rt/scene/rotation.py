"""Rotation helpers with the reference's conventions (reference src/scene/rotation.py:4-43).

Note the reference's Y rotation is the transpose of the textbook matrix (rotation.py:18-20), so a
negative pitch looks down; `euler_rotation` composes Rz(yaw) @ Ry(pitch) @ Rx(roll), degrees by default.
"""
import numpy as np


def rotation_x(phi: float) -> np.ndarray:
    c, s = np.cos(phi), np.sin(phi)
    return np.array([[1, 0, 0], [0, c, -s], [0, s, c]])


def rotation_y(theta: float) -> np.ndarray:
    c, s = np.cos(theta), np.sin(theta)
    return np.array([[c, 0, -s], [0, 1, 0], [s, 0, c]])


def rotation_z(psi: float) -> np.ndarray:
    c, s = np.cos(psi), np.sin(psi)
    return np.array([[c, -s, 0], [s, c, 0], [0, 0, 1]])


def euler_rotation(roll: float, pitch: float, yaw: float, is_radians: bool = False) -> np.ndarray:
    if not is_radians:
        roll, pitch, yaw = np.deg2rad(roll), np.deg2rad(pitch), np.deg2rad(yaw)
    return rotation_z(yaw) @ rotation_y(pitch) @ rotation_x(roll)
