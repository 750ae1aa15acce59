"""Typing aliases of the scene API (reference src/scene/common.py:1-6)."""
from typing import List, Tuple, Union
import numpy as np

Vector3D = Union[np.ndarray, list, Tuple[float, float, float]]
Color = Union[List[int], Tuple[int, int, int]]
