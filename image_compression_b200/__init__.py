"""image_compression_b200 -- B200-native JPG target-size / bilinear-downscale hot path.

Host-side mirror of the reference's seams (PolloChang/image-compression,
src/main/java/work/pollochang/compression/image/):

    ImageCompressionJpg.compressJpgWithTargetSize  core/ImageCompressionJpg.java:77
    ImageCompressionJpg.compressJpgToStream        core/ImageCompressionJpg.java:136
    ImageCompressionPng.compressPngWithTargetSize  core/ImageCompressionPng.java:37
    ImageTools.resizeImage                         tools/ImageTools.java:7

All arithmetic runs in libimgcomp_b200.so (hand-written CUDA for sm_100a) through
the C ABI in include/imgcomp_b200.h.  There is no CPU fallback: importing works
anywhere (so the symbol table can be checked without a GPU), but every compute
call raises if the library or a CUDA device is missing.
"""
from .api import (  # noqa: F401
    CompressionParams,
    Engine,
    FitResult,
    LearnedParams,
    SimilarityKey,
    ICB_OK,
    ICB_NO_FIT,
    FMT_RGB8,
    FMT_BGR8,
    FMT_RGBA8,
    FMT_ABGR8,
    FMT_GRAY8,
    FMT_ARGB32,
    FMT_XRGB32,
    IcbError,
    create_key,
    library_path,
    load_library,
    png_target_dims,
    subsample_factor,
    jpg_file_info,
    resize_dims,
)
