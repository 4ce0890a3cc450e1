// Per-image "statistics then map" operators on uint8 RGB images (whole image resident in shared memory
// in staged mode, so each image is read from HBM once and written once):
//   IMGTX_ENH_BRIGHTNESS  ImageEnhance.Brightness           (degenerate 0)
//   IMGTX_ENH_CONTRAST    ImageEnhance.Contrast             (degenerate = int(mean luma + 0.5), per image)
//                         fall_2025/tta_initial_claudeFixed.py:69-72, TTA_transforms.py:126 (ColorJitter),
//                         pipenline/cifar_image_transformations.py:81-85
//   IMGTX_ENH_COLOR       ImageEnhance.Color                (degenerate = luma per pixel)    :102-106
//   IMGTX_ENH_SHARPNESS   ImageEnhance.Sharpness            (degenerate = 3x3 SMOOTH filter) :95-99
//   IMGTX_ENH_EQUALIZE    ImageOps.equalize                 (per-channel histogram LUT) fall_2025/AugMix.py:36
//   IMGTX_ENH_HISTEQ_YUV  cv2 RGB->YUV, equalizeHist(Y), YUV->RGB  pipenline/cifar_image_transformations.py:122-129
// Roofline: HBM, 2 B per element (1 read + 1 write); the reductions run out of shared memory.
#include "enhance_op.cuh"
#include <string.h>

using namespace imgtx;

extern "C" int imgtx_enhance_u8(const uint8_t* in, uint8_t* out, int64_t n, int h, int w, int c, int kind,
                                float factor, const float* factor_per_image_dev, imgtx_stream_t stream) {
  IMGTX_CHECK_ARG(h > 0 && w > 0 && c >= 1 && c <= 4, "bad image shape");
  IMGTX_CHECK_ARG(kind >= IMGTX_ENH_BRIGHTNESS && kind <= IMGTX_ENH_HISTEQ_YUV, "unknown enhance kind %d", kind);
  if (kind == IMGTX_ENH_CONTRAST || kind == IMGTX_ENH_COLOR || kind == IMGTX_ENH_HISTEQ_YUV)
    IMGTX_CHECK_ARG(c == 3, "this enhancement needs RGB images");
  IMGTX_CHECK_ARG(c <= 3, "at most 3 channels");
  EnhanceParams p;
  p.kind = kind;
  p.factor = factor;
  p.factor_per_image = factor_per_image_dev;
  return launch_image_op<OpEnhance>(in, out, n, make_shape(h, w, c, h, w), p, (cudaStream_t)stream);
}
