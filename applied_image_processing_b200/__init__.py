"""B200-native regional colour transfer: drop-in for the hot path of the reference
``Style_3DGS/localized_style_transfer.py`` (re-exported like Style_3DGS/__init__.py:4)."""
from .localized_style_transfer import (  # noqa: F401
    LAB_TO_LMS,
    LMS_TO_LAB,
    LMS_TO_RGB,
    RGB_TO_LMS,
    RegionPCA,
    apply_pca,
    color_transfer_foreground,
    extract_foreground_deeplab,
    lab_to_rgb_image,
    lab_to_rgb_pixels,
    match_cdf,
    regional_color_transfer,
    regional_color_transfer_binned,
    regional_color_transfer_labels,
    rgb_to_lab_image,
    rgb_to_lab_pixels,
    run_localized_style_transfer,
)
from .adjacent import adain_mask_blend, deeplab_background_mask, deeplab_label_map  # noqa: F401
from .video import blend_images, blend_sequence, normalize_image, warp_blend, warp_image  # noqa: F401
from .jpeg import (  # noqa: F401
    decode_image, decode_jpeg, decode_png, encode_jpeg, image_info, jpeg_info, regional_color_transfer_files, save_jpeg,
)

__all__ = [
    "RGB_TO_LMS", "LMS_TO_LAB", "LAB_TO_LMS", "LMS_TO_RGB", "RegionPCA", "apply_pca",
    "color_transfer_foreground", "extract_foreground_deeplab", "lab_to_rgb_image", "lab_to_rgb_pixels",
    "match_cdf", "regional_color_transfer", "regional_color_transfer_binned", "regional_color_transfer_labels", "rgb_to_lab_image",
    "rgb_to_lab_pixels", "run_localized_style_transfer", "adain_mask_blend", "deeplab_background_mask",
    "deeplab_label_map", "warp_image", "normalize_image", "blend_images", "warp_blend", "blend_sequence",
]
