/*
 * orbfe.h — C ABI of the B200-native ORB feature front end (drop-in boundary).
 *
 * Plain C: pointers, sizes and POD structs only.  No OpenCV, torch or C++ types cross this line.
 * Every entry point names the reference interface it replaces (paths relative to the ORB-SLAM3
 * tree).  The C++ classes ORB_SLAM3::ORBextractor / ORBmatcher in include/ORBextractor.h and
 * include/ORBmatcher.h are thin shims over these calls, so Frame / Tracking / LocalMapping link
 * unchanged.  There is no CPU fallback: every call needs a CUDA device and fails with a negative
 * orbfe_status otherwise.
 *
 * Conventions: 0 (or a count >= 0) on success, negative orbfe_status on error; never throws, never
 * exits.  A handle is single-threaded; different handles may be used concurrently (each owns one
 * CUDA stream and its device arenas).  Matcher calls are stateless.
 */
#ifndef ORBFE_H_
#define ORBFE_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef enum orbfe_status {
    ORBFE_OK = 0,
    ORBFE_ERR_EMPTY_IMAGE = -1,   /* same value the reference returns, ORBextractor.cc:1072-1073 */
    ORBFE_ERR_BAD_ARGUMENT = -2,
    ORBFE_ERR_TOO_LARGE = -3,     /* image / batch / feature count above what the handle was created for */
    ORBFE_ERR_CAPACITY = -4,      /* caller's output buffers too small */
    ORBFE_ERR_CUDA = -5,          /* a CUDA call failed; orbfe_last_cuda_error() has the text */
    ORBFE_ERR_NO_DEVICE = -6,
    ORBFE_ERR_NOT_READY = -7      /* results requested before an extract call */
} orbfe_status;

/* Bit-compatible with cv::KeyPoint (28 bytes): the shim memcpy's into std::vector<cv::KeyPoint>. */
typedef struct orbfe_keypoint {
    float x, y;      /* pt, level-0 pixel units (ORBextractor.cc:1131-1133) */
    float size;      /* int(31 * scale[octave]) (ORBextractor.cc:862) */
    float angle;     /* degrees in [0,360), fastAtan2 of the intensity centroid (ORBextractor.cc:75-102) */
    float response;  /* FAST score */
    int32_t octave;
    int32_t class_id; /* always -1 */
} orbfe_keypoint;

/* The five scalars of ORBextractor::ORBextractor (ORBextractor.cc:408-411, parsed from yaml at
 * Tracking.cc:1022-1103) plus device placement and arena sizing. */
typedef struct orbfe_params {
    int nfeatures;
    float scale_factor;
    int nlevels;
    int ini_th_fast;
    int min_th_fast;
    int device;      /* CUDA ordinal */
    int max_width;   /* largest image the handle must accept */
    int max_height;
    int max_batch;   /* frames per extract_batch call (1 for the plain drop-in) */
} orbfe_params;

typedef struct orbfe_handle orbfe_handle;

/* ---- extractor: replaces ORB_SLAM3::ORBextractor (include/ORBextractor.h:43-109) ------------ */

int orbfe_create(const orbfe_params* params, orbfe_handle** out);
void orbfe_destroy(orbfe_handle* h);

/* Scale tables Frame copies at Frame.cc:107-113: GetScaleFactors / GetInverseScaleFactors /
 * GetScaleSigmaSquares / GetInverseScaleSigmaSquares, and the per-level quotas (ORBextractor.cc:433-444).
 * Any pointer may be NULL.  Arrays hold nlevels entries. */
int orbfe_get_tables(const orbfe_handle* h, float* scale, float* inv_scale, float* sigma2, float* inv_sigma2,
                     int* features_per_level);

/* Upper bound on keypoints per frame (sum over levels of quota + 3, ORBextractor.cc:728-729). */
int orbfe_max_keypoints(const orbfe_handle* h);

/* ORBextractor::operator()(image, mask, keypoints, descriptors, vLappingArea)  (ORBextractor.cc:1068-1150).
 * image: 8-bit single channel host buffer, `stride` bytes per row.  lap0/lap1 = vLappingArea[0..1].
 * kps/desc: caller buffers with room for `cap` keypoints (desc is cap x 32, continuous CV_8U rows).
 * *n_out = number of keypoints.  Returns monoIndex (>= 0) or a negative status; an empty image
 * returns ORBFE_ERR_EMPTY_IMAGE (-1) like the reference. */
int orbfe_extract(orbfe_handle* h, const uint8_t* image, int width, int height, ptrdiff_t stride, int lap0, int lap1,
                  orbfe_keypoint* kps, uint8_t* desc, int cap, int* n_out);

/* Batched form of the same call: `batch` frames of identical shape, frame b at image + b*frame_stride.
 * Outputs for frame b start at kps + b*cap and desc + b*cap*32; n_out[b], mono_out[b] per frame.
 * Used for stereo pairs (batch 2 replaces the two host threads of Frame.cc:119-122) and sequences. */
int orbfe_extract_batch(orbfe_handle* h, const uint8_t* images, int batch, int width, int height, ptrdiff_t stride,
                        ptrdiff_t frame_stride, int lap0, int lap1, orbfe_keypoint* kps, uint8_t* desc, int cap,
                        int* n_out, int* mono_out);

/* Same computation with the frames already in device memory and the results left in the handle's
 * device buffers (asynchronous on the handle's stream).  d_images: device pointer, same layout. */
int orbfe_extract_batch_device(orbfe_handle* h, const uint8_t* d_images, int batch, int width, int height,
                               ptrdiff_t stride, ptrdiff_t frame_stride, int lap0, int lap1);

/* Device-side result views of the last extract (valid until the next extract on this handle):
 * keypoints [batch][cap] (orbfe_keypoint), descriptors [batch][cap][32], counts [batch], mono [batch]. */
int orbfe_device_results(orbfe_handle* h, const orbfe_keypoint** d_kps, const uint8_t** d_desc, const int** d_n,
                         const int** d_mono, int* cap);
/* Copy the last results to host buffers (synchronises the handle's stream). */
int orbfe_fetch_results(orbfe_handle* h, int batch, orbfe_keypoint* kps, uint8_t* desc, int cap, int* n_out,
                        int* mono_out);

/* mvImagePyramid[level] (public member, ORBextractor.h:83; read by Frame::ComputeStereoMatches,
 * Frame.cc:801,891-906): copies the bordered level ((h+38) x (w+38), 19-px reflect-101 border) of
 * frame `frame` of the last extract into dst with dst_stride bytes per row.  dst may be NULL to
 * query the size only.  *w, *h receive the ROI size. */
int orbfe_get_pyramid_level(orbfe_handle* h, int frame, int level, uint8_t* dst, ptrdiff_t dst_stride, int* w, int* hgt);

/* Debug / parity taps of the last extract: FAST candidates of (frame, level) as int32 triples
 * (x, y, response) relative to (minBorderX, minBorderY), sorted into the reference's push order
 * (ORBextractor.cc:843-851).  Returns the count (may exceed cap; only cap triples are written). */
int orbfe_get_candidates(orbfe_handle* h, int frame, int level, int32_t* xyr, int cap);

/* The CUDA stream the handle enqueues on (a cudaStream_t), for callers that time with events. */
void* orbfe_stream(orbfe_handle* h);
int orbfe_synchronize(orbfe_handle* h);
/* Number of kernels this handle has launched since creation. */
long long orbfe_launch_count(const orbfe_handle* h);

/* Per-stage device timing (the REGISTER_TIMES switch of include/Config.h:4 / Frame.cc:116-127, done
 * with CUDA events on the handle's stream).  While enabled every extract call brackets its stages
 * with events; orbfe_get_stage_times synchronises, adds up the elapsed milliseconds per stage since
 * the last read into ms[ORBFE_NUM_STAGES], stores the number of extract calls covered in *calls and
 * resets the accumulators.  Stages: 0 pyramid, 1 grid FAST, 2 quadtree, 3 assemble, 4 orientation +
 * blur + descriptors. */
#define ORBFE_NUM_STAGES 5
int orbfe_set_profiling(orbfe_handle* h, int enabled);
int orbfe_get_stage_times(orbfe_handle* h, double* ms, long long* calls);
const char* orbfe_stage_name(int stage);

const char* orbfe_strerror(int status);
const char* orbfe_last_cuda_error(void);
/* "orbfe <version> sm_100a" */
const char* orbfe_version(void);

#ifdef __cplusplus
}
#endif
#endif /* ORBFE_H_ */
