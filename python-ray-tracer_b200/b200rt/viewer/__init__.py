from .image import convert_array_to_image, to_image_array

__all__ = ["convert_array_to_image", "to_image_array"]
