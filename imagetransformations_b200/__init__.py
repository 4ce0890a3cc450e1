"""imagetransformations_b200 — B200-native (sm_100a) implementation of the image-transformation hot path of
aaryaamoharir/ImageTransformations.

Layout: `csrc/` hand-written CUDA kernels behind the C ABI of include/imgtx.h (built into libimgtx.so),
`_lib.py` ctypes binding, `batch.py` batched device API, and one module per reference file that keeps the
reference's Python call signatures (transformation, transformations_code, AugMix, TTA_transforms,
TENT_transform, tta_initial, cifar_image_transformations, fgsm, tensor_corruptions, cifar_c, TENT_Pipelines,
Initial_Experiments, post_hoc_aaryaa).
"""
__version__ = "0.1.0"
