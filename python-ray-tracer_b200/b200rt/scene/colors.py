"""RGB colour constants, 0..255 (same names and values as reference src/scene/colors.py:1-6)."""
RED = [255, 70, 70]
GREEN = [70, 255, 70]
BLUE = [70, 70, 255]
YELLOW = [255, 255, 70]
GREY = [125, 125, 125]
MAGENTA = [139, 0, 139]
