/*
 * libimgtx — C ABI of the B200-native image-transformation hot path.
 *
 * The reference (aaryaamoharir/ImageTransformations) has no FFI of its own: its boundary is a set of
 * plain Python functions that call Pillow / OpenCV / torchvision / torch one image at a time.  Each entry
 * point below replaces the arithmetic behind one (or a family) of those functions, batched over n
 * images; the comment above each prototype cites the reference call site(s) it stands in for
 * (paths relative to the reference repository root).  INTEGRATION.md shows the ctypes binding a
 * maintainer of the reference would add.
 *
 * Conventions
 *  - Every pointer named *_dev / in / out / x / g is a DEVICE pointer owned by the caller.  The library
 *    never allocates persistent memory, never retains a pointer past the call and never frees caller
 *    memory.  uint8 images are HWC-interleaved and contiguous (n, h, w, c); float tensors are NCHW f32.
 *  - All work is enqueued on `stream` (a cudaStream_t passed as void*); no call synchronises the device.
 *  - Return value: IMGTX_OK (0) or a negative imgtx_status; imgtx_last_error() returns a thread-local
 *    message for the last failure.  No exception, abort or exit crosses this boundary.
 *  - "per-image" parameter arrays are device arrays of length n (nullable: NULL selects the scalar).
 *  - Functions ending in _host are pure host helpers (table builders); they take HOST pointers.
 */
#ifndef IMGTX_H_
#define IMGTX_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define IMGTX_VERSION 200

typedef enum {
  IMGTX_OK = 0,
  IMGTX_ERR_INVALID_ARGUMENT = -1,
  IMGTX_ERR_CUDA = -2,
  IMGTX_ERR_UNSUPPORTED = -3
} imgtx_status;

typedef void* imgtx_stream_t; /* cudaStream_t */

/* ---- library ------------------------------------------------------------------------------ */
int imgtx_version(void);
const char* imgtx_last_error(void);
/* Fills the SM count and opt-in shared memory per block of the current device. */
int imgtx_device_info(int* sm_count, int* max_smem_per_block);
/* The library keeps NO mutable process state: per-call behaviour is selected by the `flags` argument of the entry points
   that have one (0 = default).  (Development sweeps of the staged pipeline read IMGTX_* environment variables once, at the
   first launch.) */
/* Throughput mode of the CIFAR-10-C-style suite: float32 / fixed-point evaluation of the same real-valued expressions the
   default (bit-exact, float64-specified) mode evaluates; every output byte is within 1 LSB of the default mode on the same
   inputs and injected draws.  The default when the draws are made on chip (no bit-wise oracle exists for those). */
#define IMGTX_CC_FAST 1
/* Test flag of the default mode: its kernels decide most bytes with a fixed-point or float32 evaluation and replay the
   specification's literal arithmetic only near a rounding boundary; with this flag EVERY element takes the literal replay,
   so tests can check that both routes give the same bytes (also on on-chip draws). */
#define IMGTX_CC_REPLAY_ALL 2
/* `metrics_dev` of the suite entry points (nullable): device (n, 2) uint32 receiving the per-image metric of the pass —
   {sum of the output bytes, sum of (byte index + 1) * byte mod 2^32} — the small result the ranks all-gather after a pass
   (north star: "NCCL ... only to all-gather per-image outputs and metrics").  The throughput kernels compute it
   in their epilogue from the staged output; the others run imgtx_image_checksums_u8 on the output. */
int imgtx_image_checksums_u8(const uint8_t* x, int64_t n, int64_t image_bytes, uint32_t* metrics_dev, imgtx_stream_t stream);

/* ---- pointwise uint8 maps (256-entry LUT built with the library's exact f32 formula) -------- */
typedef enum {
  /* PIL.ImageEnhance.Brightness(img).enhance(param) — transformation.py:261-269,
     fall_2025/tta_initial_claudeFixed.py:64-67, pipenline/cifar_image_transformations.py:88-92 */
  IMGTX_PW_BRIGHTNESS = 0,
  /* cv2.convertScaleAbs(img, alpha=param, beta=0) — transformation.py:203-210 */
  IMGTX_PW_SCALE_ABS = 1,
  /* PIL.ImageOps.posterize(img, bits=int(param)) — fall_2025/AugMix.py:31 */
  IMGTX_PW_POSTERIZE = 2,
  /* PIL.ImageOps.solarize(img, threshold=int(param)) — fall_2025/AugMix.py:37 */
  IMGTX_PW_SOLARIZE = 3,
  /* identity copy (apply_blur(r=0), apply_rotation(0)) */
  IMGTX_PW_IDENTITY = 4
} imgtx_pointwise_op;

int imgtx_pointwise_u8(const uint8_t* in, uint8_t* out, int64_t n, int64_t image_bytes, int op,
                       float param, const float* param_per_image_dev, imgtx_stream_t stream);

/* ---- integer geometry (pure index gathers, no interpolation) -------------------------------------- */
typedef enum {
  IMGTX_GEOM_IDENTITY = 0,
  /* Image.transform(size, AFFINE, m) / Image.rotate(angle) with NEAREST resampling: m[6] are the doubles
     Pillow's Python layer hands to C (for rotate: Image.py's rounded cos/sin + centre shift).
     transformation.py:198-201, fall_2025/AugMix.py:30,32-35, fall_2025/Initial_Experiments.py:88-92 */
  IMGTX_GEOM_AFFINE = 1,
  /* apply_translation(img, dx, dy): integer shift, zero fill — transformation.py:284-307 */
  IMGTX_GEOM_SHIFT = 2,
  /* Image.transpose(FLIP_LEFT_RIGHT) — fall_2025/transformations_code:39-41,
     fall_2025/tta_initial_claudeFixed.py:52-54, T.RandomHorizontalFlip(p=1) TTA_transforms.py:75 */
  IMGTX_GEOM_FLIP_LR = 3,
  /* T.RandomCrop(out, padding=pad, padding_mode='reflect') at offset (dy=i, dx=j) — TTA_transforms.py:80 */
  IMGTX_GEOM_CROP_REFLECT = 4,
  /* exact special cases of Image.rotate */
  IMGTX_GEOM_ROT180 = 5,
  IMGTX_GEOM_ROT90 = 6,
  IMGTX_GEOM_ROT270 = 7
} imgtx_geom_kind;

typedef struct {
  int32_t kind; /* imgtx_geom_kind */
  int32_t dx, dy, pad;
  double m[6];
} imgtx_geom_t;

/* One geometry record for the whole batch (uniform_host, HOST pointer) or one per image
   (per_image_dev, DEVICE array of n records). fill_host: c bytes (NULL = zeros). */
int imgtx_geometry_u8(const uint8_t* in, uint8_t* out, int64_t n, int h, int w, int c, int oh, int ow,
                      const imgtx_geom_t* uniform_host, const imgtx_geom_t* per_image_dev,
                      const uint8_t* fill_host, imgtx_stream_t stream);

/* ---- per-image statistics + map (ImageEnhance / ImageOps.equalize / cv2 equalizeHist) ------------- */
typedef enum {
  IMGTX_ENH_BRIGHTNESS = 0, /* ImageEnhance.Brightness — transformation.py:261-269 */
  IMGTX_ENH_CONTRAST = 1,   /* ImageEnhance.Contrast — fall_2025/tta_initial_claudeFixed.py:69-72,
                               pipenline/cifar_image_transformations.py:81-85, T.ColorJitter(contrast) */
  IMGTX_ENH_COLOR = 2,      /* ImageEnhance.Color — pipenline/cifar_image_transformations.py:102-106 */
  IMGTX_ENH_SHARPNESS = 3,  /* ImageEnhance.Sharpness — pipenline/cifar_image_transformations.py:95-99 */
  IMGTX_ENH_EQUALIZE = 4,   /* ImageOps.equalize — fall_2025/AugMix.py:36 (factor ignored) */
  IMGTX_ENH_HISTEQ_YUV = 5  /* TransformationPool.histogram_equalization —
                               pipenline/cifar_image_transformations.py:122-129 (factor ignored) */
} imgtx_enhance_kind;
int imgtx_enhance_u8(const uint8_t* in, uint8_t* out, int64_t n, int h, int w, int c, int kind, float factor,
                     const float* factor_per_image_dev, imgtx_stream_t stream);

/* ---- Pillow resample (Image.resize) fused with crop-before and crop/pad-after ------------------------ */
typedef enum {
  IMGTX_FILTER_BOX = 0,
  IMGTX_FILTER_BILINEAR = 1,
  IMGTX_FILTER_BICUBIC = 2,
  IMGTX_FILTER_LANCZOS = 3
} imgtx_filter;
/* Host helpers: size (in int32 words) and contents of the Q22 coefficient table of one axis
   (libImaging/Resample.c precompute_coeffs + normalize_coeffs_8bpc).  The caller uploads the table. */
int64_t imgtx_resample_table_words(int in_size, int out_size, int filter);
int imgtx_resample_table_host(int in_size, int out_size, int filter, int32_t* table_host);
/* out canvas (oh, ow) <- zeros; r = resize(crop(in, box), (rw, rh)); canvas[place_y.., place_x..] = r
   (negative place = crop of r).  box origin uniform (crop_x, crop_y) or per image (crop_xy_per_image_dev,
   (n,2) int32 x,y).  A table pointer is NULL iff that axis keeps its size (Pillow skips the pass).
   transformation.py:173-196 (apply_scale), fall_2025/tta_initial_claudeFixed.py:56-62 (rand_crop),
   fall_2025/TTA_transforms.py:104-113 (Resize+CenterCrop) */
int imgtx_resample_u8(const uint8_t* in, uint8_t* out, int64_t n, int h, int w, int c, int crop_x, int crop_y,
                      int crop_w, int crop_h, const int32_t* crop_xy_per_image_dev, int rw, int rh,
                      const int32_t* htable_dev, int64_t htable_words, const int32_t* vtable_dev,
                      int64_t vtable_words, int oh, int ow, int place_x, int place_y, imgtx_stream_t stream);

/* Image.resize((sw, sh), f) followed by Image.resize((w, h), f) in one launch (uint8 intermediates stay in shared memory);
   tables from imgtx_resample_table_host for (w -> sw), (h -> sh), (sw -> w), (sh -> h).  CIFAR-10-C pixelate: BOX both
   ways (names only in the reference: attacks/fgsm/cifar-10/reverse_fgsm_cifar10_corrupted.py:18-23). */
int imgtx_resize_chain_u8(const uint8_t* in, uint8_t* out, int64_t n, int h, int w, int c, int sw, int sh,
                          const int32_t* h1_dev, int64_t h1_words, const int32_t* v1_dev, int64_t v1_words,
                          const int32_t* h2_dev, int64_t h2_words, const int32_t* v2_dev, int64_t v2_words,
                          imgtx_stream_t stream);

/* The same BOX down / BOX up chain for images up to 64x64 (CIFAR-10-C pixelate), throughput form: BOX up-sampling is a pure
   gather and BOX down-sampling has at most two taps, so only the small image is computed.  Returns IMGTX_OK when launched,
   1 when the shape or table structure is outside the fast path (use imgtx_resize_chain_u8: identical results). */
int imgtx_box_pixelate_u8(const uint8_t* in, uint8_t* out, int64_t n, int h, int w, int c, int sw, int sh,
                          int flags, uint32_t* metrics_dev, imgtx_stream_t stream);

/* ---- TransformationPool chains ------------------------------------------------- */
/* Individual(transformations).apply_transformations for a whole population in ONE launch
   (pipenline/cifar_image_transformations.py:141-152 over the population of :203-226): tensor_to_pil -> the ops ->
   pil_to_tensor; the uint8 image stays in shared memory between the ops. */
typedef enum {
  IMGTX_POOL_GAUSSIAN_NOISE = 0,         /* :39-48   iparam = severity 1..5 */
  IMGTX_POOL_IMPULSE_NOISE = 1,          /* :50-59   iparam = severity */
  IMGTX_POOL_SHOT_NOISE = 2,             /* :61-70   iparam = severity */
  IMGTX_POOL_DEFOCUS_BLUR = 3,           /* :72-77   iparam = severity */
  IMGTX_POOL_ENHANCE_CONTRAST = 4,       /* :81-85   fparam = factor */
  IMGTX_POOL_ENHANCE_BRIGHTNESS = 5,     /* :88-92 */
  IMGTX_POOL_ENHANCE_SHARPNESS = 6,      /* :95-99 */
  IMGTX_POOL_ENHANCE_COLOR = 7,          /* :102-106 */
  IMGTX_POOL_MOTION_BLUR = 8,            /* :109-119 iparam = size (odd) */
  IMGTX_POOL_HISTOGRAM_EQUALIZATION = 9, /* :122-129 */
  IMGTX_POOL_END = -1                    /* ends an individual's list before max_ops */
} imgtx_pool_kind;
typedef struct {
  int32_t kind;    /* imgtx_pool_kind */
  int32_t iparam;
  float fparam;
  int32_t draw;    /* noise ops: index of this op's block in its draw array; < 0 = draws made on chip from `seed` */
} imgtx_pool_op_t;
/* x: normalised f32, n images x_image_stride floats apart (0: every individual starts from the same (3,h,w) image);
   out (n,3,h,w) f32; ops_dev (n, max_ops) records.  Draw arrays (device, nullable when no op refers to them): gaussian
   blocks of h*w*3 f64 (np.random.normal(0, std*255, shape)), impulse blocks of h*w f64 (np.random.random), shot blocks of
   h*w*3 int64 (np.random.poisson).  Three copies of the uint8 image are kept on chip: sizes up to about 150x150. */
int imgtx_pool_chain_f32(const float* x, int64_t x_image_stride, float* out, int64_t n, int h, int w,
                         const imgtx_pool_op_t* ops_dev, int max_ops, const float* mean_host, const float* std_host,
                         const double* gauss_draws_dev, const double* impulse_draws_dev, const int64_t* shot_draws_dev,
                         uint64_t seed, imgtx_stream_t stream);

/* ---- stencils ------------------------------------------------------------------------------- */
/* OpenCV's bit-exact Q8 Gaussian kernel for CV_8U (ksize odd <= 31): host helper. */
int imgtx_cv_gaussian_kernel_host(int ksize, double sigma, int32_t* k_host);
/* cv2.GaussianBlur(img, (ksize, ksize), sigma) on uint8 — transformation.py:228-257 (apply_blur) */
int imgtx_cv_gaussian_blur_u8(const uint8_t* in, uint8_t* out, int64_t n, int h, int w, int c, int ksize,
                              double sigma, int flags, imgtx_stream_t stream);
/* Allocation-free variant of imgtx_cv_gaussian_blur_u8 (same bytes out).  No counterpart in the reference: added
   for callers that cannot call cudaMalloc (CUDA graph capture, tight serving loops,
   harnesses that forbid it).  The 16-bit intermediate planes live in a static device array of
   imgtx_noalloc_scratch_bytes() = 32 MiB that every GPU loading the library holds until exit; the call performs no
   cudaMalloc / cudaFree / allocating library call.  Limit: 2*h*w*c (rounded up to 256) must not exceed that size, else
   IMGTX_ERR_INVALID_ARGUMENT.  All calls on a device share the array (nothing in it needs clearing between calls), so
   calls that overlap on the GPU race: issue them on one stream. */
int64_t imgtx_noalloc_scratch_bytes(void);
int imgtx_cv_gaussian_blur_noalloc_u8(const uint8_t* in, uint8_t* out, int64_t n, int h, int w, int c, int ksize,
                                      double sigma, imgtx_stream_t stream);
/* img.filter(ImageFilter.GaussianBlur(radius)) — pipenline/cifar_image_transformations.py:72-77 */
int imgtx_pil_gaussian_blur_u8(const uint8_t* in, uint8_t* out, int64_t n, int h, int w, int c, float radius,
                               imgtx_stream_t stream);
/* cv2.filter2D with the size x size "middle row = 1/size" kernel —
   pipenline/cifar_image_transformations.py:109-119 (motion_blur) */
int imgtx_box_filter_h_u8(const uint8_t* in, uint8_t* out, int64_t n, int h, int w, int c, int size,
                          imgtx_stream_t stream);

/* ---- interpolating warps --------------------------------------------------------------------- */
/* Pillow Image.transform(size=(ow,oh), AFFINE (6 coeffs) | PERSPECTIVE (8 coeffs), resample=BILINEAR|BICUBIC,
   fillcolor): f64 evaluation identical to libImaging/Geometry.c.  coeffs_host (HOST, uniform) or
   coeffs_per_image_dev (DEVICE, (n,8) doubles).  transformation.py:212-226 (apply_shear),
   fall_2025/Initial_Experiments.py:88-92 (RandomPerspective / BILINEAR on PIL images) */
int imgtx_pil_transform_u8(const uint8_t* in, uint8_t* out, int64_t n, int h, int w, int c, int oh, int ow,
                           int perspective, int bicubic, const double* coeffs_host,
                           const double* coeffs_per_image_dev, const uint8_t* fill_host, imgtx_stream_t stream);
/* ToTensor -> torchvision RandomPerspective(p=1) on the tensor (bilinear grid_sample, fill) -> ToPILImage.
   coeffs: DEVICE (n,8) f32 from _get_perspective_coeffs.  fall_2025/transformations_code:50-66 */
int imgtx_perspective_tensor_u8(const uint8_t* in, uint8_t* out, int64_t n, int h, int w, int c,
                                const float* coeffs_per_image_dev, float fill, imgtx_stream_t stream);
/* TF.gaussian_blur(img, kernel_size, sigma) on f32 planes (reflect padding) —
   fall_2025/angellic_with_corrupt.py:54-56 */
int imgtx_gaussian_blur_f32(const float* in, float* out, int64_t planes, int h, int w, int kernel_size, float sigma,
                            imgtx_stream_t stream);

/* Entropy-gated adaptive chain on PIL images — fall_2025/Initial_Experiments.py:88-92 (the Compose) and :161-167 (its call
   site): RandomHorizontalFlip(0.5) -> RandomAffine(15, translate .1, scale .9-1.1) [Pillow AFFINE NEAREST, fill 0] ->
   RandomPerspective(.2, p=.5) [Pillow PERSPECTIVE BILINEAR, fill 0], fused into one gather per output pixel.  One record per
   image, built on the host from torchvision's own draws (imagetransformations_b200/Initial_Experiments.py). */
typedef struct {
  int32_t flags;   /* bit 0: flip fired; bit 1: affine present; bit 2: perspective fired */
  int32_t a[6];    /* libImaging affine_fixed coefficients: FIX(m0), FIX(m1), FIX(m2 + m0/2 + m1/2), FIX(m3), FIX(m4),
                      FIX(m5 + m3/2 + m4/2) with FIX(v) = floor(v*65536 + 0.5), m = the matrix torchvision hands to Pillow */
  int32_t pad;
  double persp[8]; /* _get_perspective_coeffs(startpoints, endpoints) */
} imgtx_adaptive_t;
int imgtx_adaptive_chain_u8(const uint8_t* in, uint8_t* out, int64_t n, int h, int w, int c,
                            const imgtx_adaptive_t* records_dev, imgtx_stream_t stream);

/* ---- noise ---------------------------------------------------------------------------------- */
typedef enum {
  IMGTX_NOISE_GAUSS_F32 = 0, /* transformation.py:272-281; draws: f32 (n,h,w,c) = normal(0, std*255) cast */
  IMGTX_NOISE_GAUSS_F64 = 1, /* cifar_image_transformations.py:39-48; draws: f64 (n,h,w,c) */
  IMGTX_NOISE_IMPULSE = 2,   /* :50-59; draws: f64 (n,h,w) uniform mask; param = noise_prob */
  IMGTX_NOISE_SHOT = 3       /* :61-70; draws: int64 (n,h,w,c) Poisson counts; param = lambda */
} imgtx_noise_kind;
/* parity mode: the reference's RNG draws are inputs */
int imgtx_noise_injected_u8(const uint8_t* in, uint8_t* out, int64_t n, int h, int w, int c, int kind,
                            const void* draws_dev, double param, imgtx_stream_t stream);
/* throughput mode: on-chip Philox4x32-10 keyed by (seed, element); param = std*255 | prob | lambda */
int imgtx_noise_philox_u8(const uint8_t* in, uint8_t* out, int64_t n, int h, int w, int c, int kind, double param,
                          uint64_t seed, imgtx_stream_t stream);

/* ---- CIFAR-10-C-style corruption suite ------------------------------------------------------------ */
/* The reference names the 15 corruption types (attacks/fgsm/cifar-10/reverse_fgsm_cifar10_corrupted.py:18-23) and
   consumes them as (50000,32,32,3) uint8 severity-major files (fall_2025/TTA_transforms.py:16-26,
   transformation.py:35-41); it has generator code for four of them only
   (pipenline/cifar_image_transformations.py:39-77 -> imgtx_noise_injected_u8 / imgtx_pil_gaussian_blur_u8).  The entry
   points below generate all 15 canonical types (specification: oracle/cifar_c.py).  Every random
   input is either INJECTED (device array of the draws; parity mode) or, when its pointer is NULL, drawn on chip from
   Philox4x32-10 keyed by (seed, image index, element) (throughput mode). */
typedef enum {
  IMGTX_CC_GAUSSIAN = 0,   /* p0 = sigma (on-chip mode); a = f64 (n,h,w,3) noise, already scaled */
  IMGTX_CC_SHOT = 1,       /* p0 = c; a = int64 (n,h,w,3) Poisson counts */
  IMGTX_CC_IMPULSE = 2,    /* p0 = amount; a, b = f64 (n,h,w,3) uniforms (flip, salt) */
  IMGTX_CC_SPECKLE = 3,    /* p0 = sigma; a = f64 noise */
  IMGTX_CC_BRIGHTNESS = 4, /* p0 = c (HSV value shift) */
  IMGTX_CC_CONTRAST = 5,   /* p0 = c */
  IMGTX_CC_FOG = 6,        /* p0 = c0, p1 = wibbledecay; a = f64 (n,h,w) plasma maps */
  IMGTX_CC_FROST = 7       /* p0 = c0, p1 = c1; a = uint8 (n,h,w,3) frost crops (tex_n == 0) or a texture bank
                              (tex_n, tex_h, tex_w, 3) cropped at random per image */
} imgtx_cc_kind;
int imgtx_cc_pointwise_u8(const uint8_t* in, uint8_t* out, int64_t n, int h, int w, int c, int kind, double p0, double p1,
                          const void* draws_a_dev, const void* draws_b_dev, uint64_t seed, int tex_n, int tex_h,
                          int tex_w, int flags, uint32_t* metrics_dev, imgtx_stream_t stream);
/* On-chip gaussian_noise / shot_noise sampled in the OUTPUT domain.  Input and output are bytes, so gaussian_noise adds the
   integer floor(255*sigma*Z) (one discrete law) and shot_noise is a 256-row stochastic matrix (row x = law of
   uint8(clip(Poisson(x/255*c)/c)*255)); both are sampled with Walker alias tables: one 32-bit Philox word, one multiply and
   one shared-memory lookup per element.  The table is built on the HOST in f64 (tail mass < 1e-10 dropped, thresholds
   rounded to 2^-23) by imgtx_cc_noise_table_host, uploaded by the caller and passed as table_dev (16-byte aligned).
   imgtx_cc_noise_table_words returns 0 when no table exists for the parameter (gaussian: 255*sigma > 39; shot: more than
   512 outcomes per input value); imgtx_cc_pointwise_u8 with NULL draws is the general (slower) on-chip path. */
/* kind IMGTX_CC_SHOT_FLAT (both for the table calls and for imgtx_cc_noise_onchip_u8) selects the FLAT shot_noise table:
   every row padded to the same power-of-two number K of slots, entry = threshold16 << 16 | alias output byte << 8 | own
   output byte, so an element costs ONE shared-memory lookup at word (x << log2 K) + top bits of the Philox word (no row
   record, no multiply, no output table) and the table is 256 K words (64 KB at c = 60).  Thresholds carry 16 bits: every
   outcome probability is within 2^-17 of the exact law.  Small images only (they are staged per team beside the table):
   IMGTX_ERR_UNSUPPORTED otherwise, and the caller uses the IMGTX_CC_SHOT table. */
#define IMGTX_CC_SHOT_FLAT 0x101
int64_t imgtx_cc_noise_table_words(int kind, double p0);
int imgtx_cc_noise_table_host(int kind, double p0, int32_t* table_host, int64_t capacity_words);
int imgtx_cc_noise_onchip_u8(const uint8_t* in, uint8_t* out, int64_t n, int h, int w, int c, int kind,
                             const int32_t* table_dev, int64_t table_words, uint64_t seed, int flags, uint32_t* metrics_dev, imgtx_stream_t stream);
/* defocus_blur: correlation with the sparse taps (dy, dx, weight) of the anti-aliased disk kernel, REFLECT_101, f64.
   dydx_host: (ntaps, 2) int32; |offset| <= 2. */
int imgtx_cc_defocus_u8(const uint8_t* in, uint8_t* out, int64_t n, int h, int w, int c, int ntaps,
                        const int32_t* dydx_host, const double* weights_host, int flags, uint32_t* metrics_dev, imgtx_stream_t stream);
/* zoom_blur (any fixed linear map on x/255 in CSR form, applied in f32 in stored order). */
int imgtx_cc_csr_u8(const uint8_t* in, uint8_t* out, int64_t n, int h, int w, int c, const int32_t* indptr_dev,
                    const int32_t* indices_dev, const float* weights_dev, int nnz, imgtx_stream_t stream);
/* The same operator with the caller stating the length of its longest row (max_row_taps >= max(diff(indptr)); rows are
   truncated to it): the kernel re-packs the operator into ELL form in shared memory for conflict-free access.  Falls back
   to imgtx_cc_csr_u8 when the table does not fit. */
int imgtx_cc_csr_ell_u8(const uint8_t* in, uint8_t* out, int64_t n, int h, int w, int c, const int32_t* indptr_dev,
                        const int32_t* indices_dev, const float* weights_dev, int nnz, int max_row_taps,
                        int flags, uint32_t* metrics_dev, imgtx_stream_t stream);
/* A sparse linear operator with non-negative weights summing to 1 per output pixel on (n,32,32,3) images — zoom_blur's merged
   operator (the one imgtx_cc_csr_u8 takes) — evaluated on the integer tensor cores, <= 1 LSB from imgtx_cc_csr_u8.
   tables_dev: the operator compiled by imagetransformations_b200/cifar_c.py:op32_mma_tables (Q16 weights as byte limbs in
   mma fragment order, per 4x4 output tile the aligned 4-pixel source groups it contracts over; layout in
   csrc/cifar_c_zoom.cu), table_words its length in int32, total_steps its number of MMA steps (64..192).  Replaces the
   per-image scipy zoom + mean of the canonical zoom_blur (the reference loads its output: SURVEY H27). */
int imgtx_cc_op32_mma_u8(const uint8_t* in, uint8_t* out, int64_t n, const int32_t* tables_dev, int64_t table_words,
                         int total_steps, uint32_t* metrics_dev, imgtx_stream_t stream);
/* zoom_blur, throughput form for images larger than 32x32 (<= 1 LSB from the operator form): the sum over the zoom levels
   evaluated level by level from per-axis tables (device): yidx (n_levels, 2, h) = source rows i0, i1 of every output row,
   yt (n_levels, h) = its interpolation weight, xidx / xt likewise for columns (the per-level tables the canonical
   clipped_zoom defines; the merged operator of imgtx_cc_csr_u8 is their products summed over the levels).  n_levels <= 32. */
int imgtx_cc_zoom_levels_u8(const uint8_t* in, uint8_t* out, int64_t n, int h, int w, int c, int n_levels,
                            const int32_t* yidx_dev, const float* yt_dev, const int32_t* xidx_dev, const float* xt_dev,
                            uint32_t* metrics_dev, imgtx_stream_t stream);
/* glass_blur: weights_host[0..lw] = scipy Gaussian taps (centre first); deltas_dev (n, iters, h-2, w-2, 2) int8 (dx, dy). */
int imgtx_cc_glass_u8(const uint8_t* in, uint8_t* out, int64_t n, int h, int w, int c, const double* weights_host, int lw,
                      int iters, const int8_t* deltas_dev, uint64_t seed, int flags, uint32_t* metrics_dev, imgtx_stream_t stream);
/* motion_blur: k_host[width] one-sided Gaussian line weights; offsets_dev (n, width, 2) int32 (dx, dy), or NULL with the
   angle drawn uniformly in [angle_lo, angle_hi] degrees. */
int imgtx_cc_motion_u8(const uint8_t* in, uint8_t* out, int64_t n, int h, int w, int c, int width, const float* k_host,
                       const int32_t* offsets_dev, double angle_lo, double angle_hi, uint64_t seed,
                       int flags, uint32_t* metrics_dev, imgtx_stream_t stream);
/* snow: zoom_idx_host = [y0 (h), y1 (h), x0 (w), x1 (w)], zoom_t_host = [ty (h), tx (w)] (centre-zoom tables of the noise
   layer); layer_dev (n,h,w) f64 normal(loc, scale) draws or NULL. */
int imgtx_cc_snow_u8(const uint8_t* in, uint8_t* out, int64_t n, int h, int w, int c, float thresh, float c6,
                     float one_minus_c6, const int32_t* zoom_idx_host, const float* zoom_t_host, int width,
                     const float* k_host, const int32_t* offsets_dev, const double* layer_dev, double loc, double scale,
                     double angle_lo, double angle_hi, uint64_t seed, int flags, uint32_t* metrics_dev, imgtx_stream_t stream);
/* elastic_transform: M_dev (n,6) f64 forward affine matrices (cv2.getAffineTransform) or NULL (corner jitter of amplitude
   c2 drawn on chip); ux_dev / uy_dev (n,h,w) f64 uniform(-1,1) fields or NULL; weights_host[0..lw] Gaussian taps (lw = -1:
   no smoothing). */
int imgtx_cc_elastic_u8(const uint8_t* in, uint8_t* out, int64_t n, int h, int w, int c, double alpha,
                        const double* weights_host, int lw, const double* M_dev, const double* ux_dev,
                        const double* uy_dev, double c2, uint64_t seed, int flags, uint32_t* metrics_dev, imgtx_stream_t stream);
/* jpeg_compression: baseline round trip (4:2:0, islow DCT) with the two quantisation tables (natural order). */
int imgtx_cc_jpeg_u8(const uint8_t* in, uint8_t* out, int64_t n, int h, int w, int c, const int32_t* qtables_host,
                     int flags, uint32_t* metrics_dev, imgtx_stream_t stream);

/* ---- tensors ------------------------------------------------------------------------------ */
/* T.ToTensor() [+ T.Normalize(mean, std) when mean_host/std_host != NULL]: u8 NHWC -> f32 NCHW.
   fall_2025/TENT_transform.py:56-59, fall_2025/TTA_transforms.py:62-65,
   fall_2025/tta_initial_claudeFixed.py:43-46, attacks/normal_inference/cifar_corrupted.py:42-48 */
int imgtx_to_tensor_f32(const uint8_t* in, float* out, int64_t n, int h, int w, int c, const float* mean_host,
                        const float* std_host, imgtx_stream_t stream);

#define IMGTX_TOU8_DENORMALIZE 1 /* x*std + mean first */
#define IMGTX_TOU8_CLAMP01 2     /* torch.clamp(x, 0, 1) */
/* tensor_to_pil (DENORMALIZE|CLAMP01) / TF.to_pil_image (flags 0: truncation with mod-256 wrap):
   f32 NCHW -> u8 NHWC.  fall_2025/tta_initial_claudeFixed.py:37-41,
   pipenline/cifar_image_transformations.py:27-30, fall_2025/AugMix.py:56 */
int imgtx_to_uint8_f32(const float* in, uint8_t* out, int64_t n, int h, int w, int c, const float* mean_host,
                       const float* std_host, int flags, imgtx_stream_t stream);

/* ---- fused AugMix ---------------------------------------------------------------------------- */
typedef enum {
  IMGTX_AUG_NONE = 0,      /* chain shorter than `depth`, or a rotate by 0 degrees */
  IMGTX_AUG_GEOM = 1,      /* Pillow NEAREST affine: a[0..5] = 16.16 fixed-point (a0,a1,a2,a3,a4,a5);
                              rotate / shear_x / shear_y / translate_x / translate_y — fall_2025/AugMix.py:30,32-35 */
  IMGTX_AUG_POSTERIZE = 2, /* param = byte mask — AugMix.py:31 */
  IMGTX_AUG_SOLARIZE = 3,  /* param = threshold — AugMix.py:37 */
  IMGTX_AUG_EQUALIZE = 4   /* ImageOps.equalize — AugMix.py:36 */
} imgtx_augmix_kind;
typedef struct {
  int32_t kind;
  int32_t a[6];
  int32_t param;
} imgtx_augmix_op_t;
/* augmix(image_tensor, severity, width, depth) — fall_2025/AugMix.py:45-62 — for n images in one launch.
   x/out: (n,3,h,w) f32.  ops_dev: (n, width, depth) records (op order within a chain = application order).
   weights_dev: (n, width+2) f32 = [f32(ws_0..ws_{width-1}), f32(m), f32(1-m)] (1-m evaluated in double first). */
int imgtx_augmix_f32(const float* x, float* out, int64_t n, int h, int w, int width, int depth,
                     const imgtx_augmix_op_t* ops_dev, const float* weights_dev, imgtx_stream_t stream);

/* ---- fused TTA view expansion --------------------------------------------------------------------- */
typedef enum {
  IMGTX_VIEW_IDENTITY = 0,           /* ToTensor+Normalize only */
  IMGTX_VIEW_FLIP = 1,               /* FLIP_LEFT_RIGHT / RandomHorizontalFlip(p=1) */
  IMGTX_VIEW_CROP_RESIZE = 2,        /* crop (p0,p1,p0+i0,p1+i0) then Image.resize((w,h)) — rand_crop; tables: i0->w, i0->h */
  IMGTX_VIEW_BRIGHTNESS = 3,         /* ImageEnhance.Brightness(p0) */
  IMGTX_VIEW_CONTRAST = 4,           /* ImageEnhance.Contrast(p0) */
  IMGTX_VIEW_CROP_REFLECT = 5,       /* RandomCrop(size, padding=i0, 'reflect') at row p0, col p1 */
  IMGTX_VIEW_FLIP_CROP_REFLECT = 6,  /* flip, then the reflect crop */
  IMGTX_VIEW_RESIZE_CENTER_CROP = 7  /* Resize(s) + CenterCrop: tables w->s, h->s; window origin (i0 left, i1 top) */
} imgtx_view_kind;
typedef struct {
  int32_t kind;
  int32_t table_h, table_v; /* indices into the table blob (resample views only) */
  int32_t i0, i1;
  int32_t reserved[3];
} imgtx_view_t;
/* One image -> n_views normalised f32 CHW views, (n, n_views, 3, h, w), in one pass.
   in: uint8 (n,h,w,3), or (in_is_f32) the normalised f32 (n,3,h,w) tensor, quantised first exactly like
   tensor_to_pil(x*in_std + in_mean, clamp, *255 trunc).  views_host: HOST array, uniform over the batch;
   params_dev: DEVICE (n, n_views, 2) f32 per-image parameters (p0, p1); tables_dev: DEVICE blob of Q22 tables from
   imgtx_resample_table_host, table_offsets_host[n_tables+1] word offsets.
   fall_2025/tta_initial_claudeFixed.py:37-78,218-230; fall_2025/TTA_transforms.py:69-138 */
int imgtx_views_f32(const void* in, int in_is_f32, float* out, int64_t n, int h, int w, int n_views,
                    const imgtx_view_t* views_host, const float* params_dev, const int32_t* tables_dev,
                    const int32_t* table_offsets_host, int n_tables, const float* in_mean_host,
                    const float* in_std_host, const float* out_mean_host, const float* out_std_host,
                    imgtx_stream_t stream);

/* ---- sign-step family --------------------------------------------------------------------- */
#define IMGTX_MAX_EPS 8
typedef enum { IMGTX_CLAMP_NONE = 0, IMGTX_CLAMP_SCALAR = 1, IMGTX_CLAMP_PER_CHANNEL = 2 } imgtx_clamp_mode;

/* out[e] = clamp(x + direction * eps[e] * sign(g)) for e < n_eps, x/g (n,c,hw) f32, out (n_eps,n,c,hw);
   pert (nullable, same shape as out) receives eps[e]*sign(g).
   direction=+1, CLAMP_NONE: FGSM            attacks/fgsm.py:190-192, :250-251
   direction=+1, CLAMP_SCALAR [0,1]: fgsm_attack   attacks/fgsm/cifar-10/fgsm_cifar10.py:25-28
   direction=-1, CLAMP_PER_CHANNEL: reverse_fgsm_attack  attacks/fgsm/cifar-10/reverse_fgsm_cifar10.py:74-80
   direction=-1, CLAMP_SCALAR: patch / entropy steps  fall_2025/angellic_patches_fgsm.py:67-69,
                                                       pipenline/cifar100_shannon.py:72-74 */
int imgtx_sign_step_f32(const float* x, const float* g, float* out, float* pert, int64_t n, int c, int64_t hw,
                        const float* eps_host, int n_eps, float direction, int clamp_mode,
                        const float* lo_host, const float* hi_host, imgtx_stream_t stream);

/* BASELINE config 5: the sign step fused with the post-attack leg of pipenline/cifar_image_transformations.py:141-152 —
   tensor_to_pil (x*std + mean, clamp [0,1], x255, truncate; :27-30) -> one pointwise uint8 op given as a 256-entry map
   (lut_host; e.g. the ImageEnhance.Brightness map of :88-92) -> pil_to_tensor ((b/255 - mean)/std; :32-34).  x and g are
   read once; out is (n_eps, n, c, hw) f32.  Same step / clamp arguments as imgtx_sign_step_f32. */
int imgtx_sign_step_post_f32(const float* x, const float* g, float* out, int64_t n, int c, int64_t hw,
                             const float* eps_host, int n_eps, float direction, int clamp_mode, const float* lo_host,
                             const float* hi_host, const float* mean_host, const float* std_host,
                             const uint8_t* lut_host, imgtx_stream_t stream);

/* PGD iteration: xa = x + alpha*sign(g); eta = clamp(xa - x0, -eps, eps); out = clamp(x0 + eta, lo, hi)
   attacks/fgsm/cifar-10/pgd_cifar10.py:37-45, pgd_cifar10_corrupted.py:72-83 */
int imgtx_pgd_step_f32(const float* x, const float* x0, const float* g, float* out, int64_t n, int c,
                       int64_t hw, float alpha, float eps, int clamp_mode, const float* lo_host,
                       const float* hi_host, imgtx_stream_t stream);

/* normalize_tensor / denormalize_tensor helpers on (n,c,hw) f32: denormalize == 0: (x - mean[ch]) / std[ch];
   != 0: x * std[ch] + mean[ch]; two rounded f32 operations per element as torch evaluates them
   (pipenline/cifar_image_transformations.py:19-25, fall_2025/AugMix.py:25-34,
   fall_2025/tta_initial_claudeFixed.py:27-35). */
int imgtx_channel_affine_f32(const float* x, float* out, int64_t n, int c, int64_t hw, const float* mean_host,
                             const float* std_host, int denormalize, imgtx_stream_t stream);

typedef enum {
  IMGTX_FC_GAUSSIAN_NOISE = 0, /* (x + (r*a + b)).clamp(0,1)        pipenline/angellic.py:34-38 */
  IMGTX_FC_BRIGHTNESS = 1,     /* (x + a).clamp(0,1)                pipenline/angellic.py:40-42 */
  IMGTX_FC_CONTRAST = 2,       /* ((x-0.5)*a + 0.5).clamp(0,1)      pipenline/angellic.py:44-46 */
  IMGTX_FC_ADD_NOISE = 3       /* x + r*a   (no clamp)              fall_2025/angellic_with_corrupt.py:51-52 */
} imgtx_float_corrupt_mode;
/* noise: injected standard-normal draws (torch.randn_like), same shape as x; NULL for modes 1,2. */
int imgtx_float_corrupt_f32(const float* x, const float* noise, float* out, int64_t count, int mode, float a,
                            float b, int do_clamp, imgtx_stream_t stream);

/* ---- per-image metric: Shannon-entropy gate -------------------------------- */
/* compute_shannon_entropy for a batch: x f32 (n,c,h,w) -> entropy_dev[n] (f64, bits) of the 256-bin histogram over all
   c*h*w values; counts_dev (optional, n*256 u32) receives the integer histogram np.histogram(bins=256, range=(0,1))
   would give.  mean_host/std_host != NULL: the tensor is normalised and is first mapped x*std+mean (f32) and clamped
   to [0,1] (fall_2025/TENT_Pipelines.py:99-121, caller :257-274); NULL: values are used as they are and those
   outside [0,1] are not counted (fall_2025/Initial_Experiments.py:95-113, caller :147-157).  The counts are exact;
   the entropy is summed in f64 in a fixed order (within 1e-12 of SciPy's value).  An image with no countable value
   gives NaN, as the reference does. */
int imgtx_shannon_entropy_f32(const float* x, int64_t n, int c, int h, int w, const float* mean_host,
                              const float* std_host, double* entropy_dev, uint32_t* counts_dev, imgtx_stream_t stream);

/* ---- tensor-backend TTA chain ------------------------------------------------ */
/* One record per OUTPUT image of TestTimeAugmentation (fall_2025/post_hoc_aaryaa.py:24-51): the draws of
   Compose([RandomHorizontalFlip, RandomRotation(10), RandomAffine(0, scale), ColorJitter(brightness),
   RandomPerspective]) already turned into torchvision's parameters on the host. */
typedef struct {
  int32_t flip;        /* RandomHorizontalFlip fired */
  int32_t perspective; /* RandomPerspective fired */
  float brightness;    /* ColorJitter brightness factor (f32 cast of the Python float) */
  float reserved;
  float rot[6];        /* _get_inverse_affine_matrix([0,0], -angle, [0,0], 1, [0,0]) cast to f32 */
  float aff[6];        /* _get_inverse_affine_matrix([0,0], 0, [0,0], scale, [0,0]) cast to f32 */
  float persp[8];      /* _get_perspective_coeffs(startpoints, endpoints) cast to f32 */
} imgtx_tta_chain_t;
/* x f32 (n,c,h,w) -> out f32 (n*views,c,h,w): output image j is the chain chains_dev[j] applied to input image
   j / views.  NEAREST rotate/affine through torchvision's f32 grid (bit-identical picks), brightness blend + clamp,
   BILINEAR perspective with the ones-mask fill (within 1e-5*max(1,|v|) of torch's grid_sample: summation order). */
int imgtx_tensor_tta_f32(const float* x, float* out, int64_t n, int views, int c, int h, int w,
                         const imgtx_tta_chain_t* chains_dev, imgtx_stream_t stream);
/* Mean over views: in f32 (views, count) -> out[i] = (((0 + in[0][i]) + in[1][i]) + ...) / views
   (fall_2025/post_hoc_aaryaa.py:334-342 mean logits; fall_2025/TTA_transforms.py:208-209 mean probabilities). */
int imgtx_mean_views_f32(const float* in, float* out, int64_t count, int views, imgtx_stream_t stream);

#ifdef __cplusplus
}
#endif
#endif /* IMGTX_H_ */
