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

/* The whole mvImagePyramid of one frame of the last extract with ONE synchronise: dst[l] receives the bordered level l
 * ((h_l + 38) rows of w_l + 38 bytes, dst_stride[l] bytes apart); sizes from orbfe_get_level_size (valid once an image of
 * the current shape has been extracted).  What the C++ shim calls after operator() so that unmodified host readers of
 * ORBextractor::mvImagePyramid (Frame::ComputeStereoMatches, Frame.cc:801,891-906) keep working. */
int orbfe_get_pyramid(orbfe_handle* h, int frame, uint8_t* const* dst, const ptrdiff_t* dst_stride);
int orbfe_get_level_size(const orbfe_handle* h, int level, int* w, int* hgt);

/* Debug / parity taps of the last extract: FAST candidates of (frame, level) as int32 triples
 * (x, y, response) relative to (minBorderX, minBorderY), sorted into the reference's push order
 * (ORBextractor.cc:843-851).  Returns the count (may exceed cap; only cap triples are written). */
int orbfe_get_candidates(orbfe_handle* h, int frame, int level, int32_t* xyr, int cap);

/* The handle's public CUDA stream (a cudaStream_t): every call is ordered on it — internally an extraction forks
 * into side compute streams and joins back — so events recorded on it bracket whole calls. */
void* orbfe_stream(orbfe_handle* h);
int orbfe_synchronize(orbfe_handle* h);
/* Number of kernels this handle has launched since creation (kernels inside a replayed CUDA graph included), and the
 * number of cudaGraphLaunch calls among them. */
long long orbfe_launch_count(const orbfe_handle* h);
long long orbfe_graph_launch_count(const orbfe_handle* h);

/* Ordering against streams the handle does not own.  Matchers and vocabularies enqueue on their own streams; a
 * consumer that reads an extractor's DEVICE results without a host synchronise in between must be ordered after
 * the extraction, and the next extraction (which overwrites those results and the pyramid) after the consumer:
 *   orbfe_stream_wait_extractor(h, s)  work enqueued on stream `s` from now on waits for everything enqueued on h so far;
 *   orbfe_extractor_wait_stream(h, s)  work enqueued on h from now on waits for everything enqueued on `s` so far.
 * The entry points that take an orbfe_handle* themselves (orbfe_stereo_match*) do both internally. */
int orbfe_stream_wait_extractor(orbfe_handle* h, void* stream);
int orbfe_extractor_wait_stream(orbfe_handle* h, void* stream);

/* Tuning of the device-side pipeline (defaults: chunks of up to 512 frames over 2 compute streams, CUDA graphs on).
 * chunk_frames / compute_streams <= 0 and use_graphs < 0 keep the current value.  Synchronises the handle. */
int orbfe_set_pipeline(orbfe_handle* h, int chunk_frames, int compute_streams, int use_graphs);

/* Which kernel computes GaussianBlur(7x7, 2, 2) (ORBextractor.cc:1114-1115): the integer-ALU kernel (k_blur) or the tensor-core
 * kernel (k_blur_tc: the two passes as banded int8 GEMMs on tcgen05.mma).  Same bytes either way; no counterpart in the
 * reference.  ORBFE_BLUR=alu|tensor in the environment sets the initial value.  Synchronises the handle. */
enum { ORBFE_BLUR_ALU = 0, ORBFE_BLUR_TENSOR = 1 };
int orbfe_set_blur_backend(orbfe_handle* h, int backend);

/* Tap: the blurred level the descriptors of the last call were sampled from (the reference's `workingMat` after GaussianBlur),
 * ROI only, w x h bytes (sizes from orbfe_get_pyramid_level) into dst with dst_stride >= w. */
int orbfe_get_blurred_level(orbfe_handle* h, int frame, int level, uint8_t* dst, ptrdiff_t dst_stride);

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

/* ---- matchers: replace ORB_SLAM3::ORBmatcher (include/ORBmatcher.h:35-111) and the two stereo ---
 * ---- matchers of Frame (Frame.cc:794-964, 1116-1156) -------------------------------------------- */

/* ORBmatcher::TH_LOW / TH_HIGH / HISTO_LENGTH (ORBmatcher.cc:36-38). */
#define ORBFE_TH_LOW 50
#define ORBFE_TH_HIGH 100
#define ORBFE_HISTO_LENGTH 30

/* ORBmatcher::DescriptorDistance (ORBmatcher.cc:2561-2577): Hamming distance of two 32-byte
 * descriptors.  A static host function in the reference (one pair, host cv::Mat rows), so it stays
 * a host function here; every bulk distance computation goes through the device entry points. */
static inline int orbfe_descriptor_distance(const uint8_t* a, const uint8_t* b) {
    int d = 0, i;
    for (i = 0; i < 32; i += 4) {
        uint32_t x = ((uint32_t)(a[i] ^ b[i])) | ((uint32_t)(a[i + 1] ^ b[i + 1]) << 8) |
                     ((uint32_t)(a[i + 2] ^ b[i + 2]) << 16) | ((uint32_t)(a[i + 3] ^ b[i + 3]) << 24);
        d += __builtin_popcount(x);
    }
    return d;
}

/* An ORBmatcher(nnratio, checkOri) object (ORBmatcher.h:39): owns one CUDA stream, pinned staging
 * and growable device scratch.  Single-threaded per handle; handles are independent. */
typedef struct orbfe_matcher orbfe_matcher;
int orbfe_matcher_create(float nnratio, int check_orientation, int device, orbfe_matcher** out);
void orbfe_matcher_destroy(orbfe_matcher* m);
void* orbfe_matcher_stream(orbfe_matcher* m);
int orbfe_matcher_synchronize(orbfe_matcher* m);
long long orbfe_matcher_launch_count(const orbfe_matcher* m);

/* DBoW2::FeatureVector (Thirdparty/DBoW2/DBoW2/FeatureVector.h:24-25: std::map<NodeId,
 * std::vector<unsigned>>) flattened to CSR: node_ids ascending, the features of node k are
 * indices[offsets[k] .. offsets[k+1]), ascending as DBoW2 appends them. */
typedef struct orbfe_feature_vector {
    int nnodes;
    const uint32_t* node_ids; /* [nnodes] */
    const int32_t* offsets;   /* [nnodes + 1] */
    const int32_t* indices;   /* [offsets[nnodes]] */
} orbfe_feature_vector;

/* cv::BFMatcher(NORM_HAMMING).knnMatch(q, t, k=2) (Frame.cc:43,1134): the two nearest train rows
 * of every query row, ties to the lowest train index.  idx2 / dist2: nq x 2 int32, -1 where the
 * train set has fewer than two rows.  Host buffers. */
int orbfe_hamming_knn2(orbfe_matcher* m, const uint8_t* q, int nq, const uint8_t* t, int nt, int32_t* idx2,
                       int32_t* dist2);
/* Batched device-resident form: problem b matches the d_nq[b] rows at d_q + b*q_stride against the
 * d_nt[b] rows at d_t + b*t_stride (d_nq / d_nt: device int arrays, or NULL for nq_max / nt_max
 * everywhere).  Results into device arrays d_idx2 / d_dist2 of [batch][nq_max][2].  Asynchronous on
 * the matcher's stream.  Used for sequence matching frame t <-> t+1 straight from an extractor's
 * device results. */
int orbfe_hamming_knn2_batch_device(orbfe_matcher* m, const uint8_t* d_q, ptrdiff_t q_stride, const int* d_nq, int nq_max,
                                    const uint8_t* d_t, ptrdiff_t t_stride, const int* d_nt, int nt_max, int batch,
                                    int32_t* d_idx2, int32_t* d_dist2);
/* The brute-force 2-NN has two exact implementations: XOR + POPC on the integer pipe, and the tensor cores (tcgen05.mma
 * kind::i8 on descriptors expanded to signed bytes; the accumulator is 256 * hamming + index).  Both give
 * identical indices and distances.  Default: ORBFE_KNN2_DEFAULT, overridden per process by the environment variable
 * ORBFE_KNN2=popc|tensor and per matcher by this call. */
enum { ORBFE_KNN2_POPC = 0, ORBFE_KNN2_TENSOR = 1 };
#define ORBFE_KNN2_DEFAULT ORBFE_KNN2_TENSOR
int orbfe_matcher_set_knn2_backend(orbfe_matcher* m, int backend);

/* Allocation-free form of orbfe_hamming_knn2 on DEVICE buffers.  No counterpart in the reference: added
 * for callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops,
 * harnesses that forbid it).  Needs no orbfe_matcher; calls no cudaMalloc / cudaFree / allocating routine.  The kernels'
 * scratch (per-split partial results) lives in a static device array of ORBFE_NOALLOC_SCRATCH_BYTES that every GPU loading
 * the library holds until exit; hence the limits nq <= ORBFE_NOALLOC_MAX_QUERIES and nt <= ORBFE_NOALLOC_MAX_TRAIN
 * (larger sizes return ORBFE_ERR_TOO_LARGE, text in orbfe_last_cuda_error()).  d_q (nq x 32), d_t (nt x 32): device
 * descriptors; d_idx2 / d_dist2: device int32 [nq][2], same contents as orbfe_hamming_knn2.  Asynchronous on `stream`
 * (a cudaStream_t of `device`, NULL = the default stream).  All calls share the one scratch array: calls that may overlap
 * on the GPU (different streams without ordering between them) race; calls on one stream are safe. */
#define ORBFE_NOALLOC_MAX_QUERIES 8192
#define ORBFE_NOALLOC_MAX_TRAIN 65536
#define ORBFE_NOALLOC_SCRATCH_BYTES (1u << 20)
int orbfe_hamming_knn2_noalloc(int device, void* stream, const uint8_t* d_q, int nq, const uint8_t* d_t, int nt, int32_t* d_idx2,
                               int32_t* d_dist2);
/* Frame::ComputeStereoFishEyeMatches up to the descriptor test (Frame.cc:1134-1141): 2-NN + Lowe
 * ratio (d0 < d1 * ratio in double precision, ratio = the double literal 0.7 there).  match[nq] = train index or -1.  Returns the number
 * of matches. */
int orbfe_match_lapping(orbfe_matcher* m, const uint8_t* q, int nq, const uint8_t* t, int nt, double ratio,
                        int32_t* match);

/* Frame::ComputeStereoFishEyeMatches (Frame.cc:1116-1156), whole function: 2-NN + Lowe ratio 0.7 between
 * the lapping-area features (rows mono_left.. of the left list, mono_right.. of the right list), then
 * KannalaBrandt8::TriangulateMatches (KannalaBrandt8.cpp:338-407: parallax cos < 0.9998, both depths
 * positive, reprojection error^2 <= 5.991 sigma^2 in both views) and depth > 1e-4.
 * cam1 / cam2: fx, fy, cx, cy, k1..k4 of the two Kannala-Brandt cameras; R12 (row-major 3x3), t12:
 * mRlr / mtlr; sigma2: mvLevelSigma2[nlevels].  Outputs (host): left_to_right[n_left],
 * right_to_left[n_right] (-1 none), depth[n_left] (-1 none), p3d[n_left * 3] (mvStereo3Dpoints).
 * The triangulation ends in a 4x4 SVD (cv::SVD in the reference): results agree to float tolerance.
 * Returns nMatches or a negative status. */
int orbfe_fisheye_stereo_match(orbfe_matcher* m, const orbfe_keypoint* kps_left, const uint8_t* desc_left, int n_left,
                               int mono_left, const orbfe_keypoint* kps_right, const uint8_t* desc_right, int n_right,
                               int mono_right, const float* cam1, const float* cam2, const float* R12, const float* t12,
                               const float* sigma2, int nlevels, int32_t* left_to_right, int32_t* right_to_left, float* depth,
                               float* p3d);

/* Colour input.  Tracking::GrabImageMonocular / Stereo / RGBD convert their input with
 * cv::cvtColor(RGB2GRAY | BGR2GRAY | RGBA2GRAY | BGRA2GRAY) before extraction (Tracking.cc:1233-1258,
 * 1325-1338); with a colour format set, orbfe_extract* take the interleaved colour image (stride in bytes) and the
 * conversion — OpenCV's 8-bit fixed point, (R * 9798 + G * 19235 + B * 3735 + 16384) >> 15 — runs inside the
 * pyramid's level-0 kernel, so no grey image is ever stored and level 0 of the pyramid equals mImGray. */
enum { ORBFE_FORMAT_GRAY8 = 0, ORBFE_FORMAT_RGB8 = 1, ORBFE_FORMAT_BGR8 = 2, ORBFE_FORMAT_RGBA8 = 3, ORBFE_FORMAT_BGRA8 = 4 };
int orbfe_set_input_format(orbfe_handle* h, int format);

/* Frame::ComputeStereoMatches (Frame.cc:794-964), "SearchStereo": rectified-stereo row-band search
 * (Hamming < 75 among right keypoints whose row band covers the left keypoint, octave within +-1,
 * disparity in [0, bf/b]), 11x11 SAD refinement over +-5 px on the pyramid level of the left
 * keypoint, parabola sub-pixel fit, 2.1 x median SAD outlier cut.  Works on the DEVICE-resident
 * results and pyramids of the last extract of `left` (frame index frame_l) and `right` (frame
 * index frame_r); the two handles may be the same object (batch extract of the pair) and must live
 * on one device.  u_right / depth: host arrays of n_left floats (mvuRight / mvDepth, -1 = none).
 * Returns the number of left keypoints with a depth. */
int orbfe_stereo_match(orbfe_matcher* m, orbfe_handle* left, int frame_l, orbfe_handle* right, int frame_r, float bf,
                       float b, float* u_right, float* depth, int cap);
/* Batched device-resident form for stereo sequences: pair p = frames (2p, 2p+1) of one batch
 * extract on `h`.  Results stay on the device: d_u_right / d_depth are [npairs][cap] floats
 * (cap = orbfe_max_keypoints(h)).  Asynchronous on the matcher's stream (ordered after the
 * extractor's stream). */
int orbfe_stereo_match_batch_device(orbfe_matcher* m, orbfe_handle* h, int npairs, float bf, float b, float* d_u_right,
                                    float* d_depth);

/* One side of a BoW-guided search. */
typedef struct orbfe_bow_side {
    const uint8_t* desc;          /* n x 32 */
    int n;
    orbfe_feature_vector fv;
    const uint8_t* mask;          /* n bytes; meaning depends on the call, may be NULL where stated */
    const orbfe_keypoint* kps;    /* n keypoints (angle, pt, octave are read); may be NULL if unused */
} orbfe_bow_side;

/* ORBmatcher::SearchByBoW, both overloads, `nprob` independent problems in one call (Tracking::
 * Relocalization and LoopClosing call it once per candidate keyframe).
 *   form 0  SearchByBoW(KeyFrame*, Frame&, vpMapPointMatches)   ORBmatcher.cc:269-471
 *           A = keyframe (mask = has a good MapPoint), B = frame (mask NULL); accept bestDist1 <= TH_LOW;
 *           nleft_b = F.Nleft (-1 without a second camera) selects the left/right split of :330-368.
 *           Output b2a[p][B.n]: keyframe feature matched to each frame feature, -1 none.
 *   form 1  SearchByBoW(KeyFrame*, KeyFrame*, vpMatches12)      ORBmatcher.cc:823-963
 *           A = pKF1, B = pKF2, both masks = has a good MapPoint; accept bestDist1 < TH_LOW.
 *           Output a2b[p][A.n]: pKF2 feature matched to each pKF1 feature, -1 none.
 * Both outputs are always written (either may be NULL).  If b_shared != 0 every problem uses B[0].
 * nmatches[p] = the function's return value.  Returns 0 or a negative status. */
int orbfe_search_bow(orbfe_matcher* m, int form, const orbfe_bow_side* A, const orbfe_bow_side* B, int nprob, int b_shared,
                     int nleft_b, int32_t* const* a2b, int32_t* const* b2a, int* nmatches);

/* Two-view geometry of SearchForTriangulation for pinhole cameras. */
typedef struct orbfe_epipolar {
    float F12[9];      /* row-major K1^-T [t12]x R12 K2^-1, as Pinhole::epipolarConstrain_ forms it (Pinhole.cpp:161-164) */
    float ex, ey;      /* projection of camera 1's centre into image 2 (ORBmatcher.cc:1213-1218) */
    const float* sigma2_b;  /* pKF2->mvLevelSigma2 [nlevels] */
    const float* scale_b;   /* pKF2->mvScaleFactors [nlevels] */
    int nlevels;
} orbfe_epipolar;

/* ORBmatcher::SearchForTriangulation / SearchForTriangulation_ (ORBmatcher.cc:965-1432), pinhole
 * cameras without a second camera.  A = pKF1, B = pKF2; mask = feature already HAS a MapPoint
 * (skipped); u_right_* = mvuRight (>= 0: stereo observation; NULL = all mono).  kps = mvKeysUn.
 * matches12[A.n] = pKF2 index or -1; the reference's vMatchedPairs is the ascending list of
 * (i, matches12[i]).  Returns the match count or a negative status. */
int orbfe_search_triangulation(orbfe_matcher* m, const orbfe_bow_side* A, const float* u_right_a, const orbfe_bow_side* B,
                               const float* u_right_b, const orbfe_epipolar* geo, int only_stereo, int coarse,
                               int32_t* matches12);

/* Two-view geometry of SearchForTriangulation for Kannala-Brandt (fisheye) keyframes. */
typedef struct orbfe_kb8_epipolar {
    float cam_a[2][8];  /* pKF1: [0] = mpCamera, [1] = mpCamera2 (ignored when nleft_a < 0); fx, fy, cx, cy, k1..k4 */
    float cam_b[2][8];  /* pKF2 likewise */
    int nleft_a;        /* pKF1->NLeft: features [0, nleft) belong to mpCamera, the rest to mpCamera2; -1 = one camera */
    int nleft_b;
    float R12[4][9];    /* x1 = R12 x2 + t12 for the pairings ll, lr, rl, rr (ORBmatcher.cc:1248-1262), row-major; */
    float t12[4][3];    /* entry 0 alone is used unless both keyframes have two cameras (:1332) */
    float ex, ey;       /* epipole in image 2; tested (:1321-1327) only when pKF1 has one camera */
    const float* sigma2_a;  /* pKF1->mvLevelSigma2 [nlevels] */
    const float* sigma2_b;  /* pKF2->mvLevelSigma2 [nlevels] */
    const float* scale_b;   /* pKF2->mvScaleFactors [nlevels] */
    int nlevels;
} orbfe_kb8_epipolar;

/* ORBmatcher::SearchForTriangulation / SearchForTriangulation_ (ORBmatcher.cc:965-1432) for keyframes with
 * KannalaBrandt8 cameras: the geometric test is KannalaBrandt8::epipolarConstrain_ (KannalaBrandt8.cpp:239-242,
 * i.e. TriangulateMatches_ (:409-480) returns a depth > 1e-4) with the cameras and relative pose of the
 * pairing (left/right camera of each keyframe) the two features belong to.  A.kps / B.kps: mvKeys followed
 * by mvKeysRight.  Fisheye keyframes carry no mvuRight, hence only_stereo != 0 matches nothing (:1280-1283).
 * matches12[A.n] as orbfe_search_triangulation.  points12 (nullable, A.n x 3 floats): the triangulated point
 * in camera-1 coordinates of every row with matches12 >= 0 (what the matchAndtriangulate overload,
 * :1435-1624, hands back, there in world coordinates); zero with coarse != 0.  The triangulation ends in a
 * 4x4 SVD: decisions agree with the reference except within float rounding of a threshold. */
int orbfe_search_triangulation_kb8(orbfe_matcher* m, const orbfe_bow_side* A, const orbfe_bow_side* B,
                                   const orbfe_kb8_epipolar* geo, int only_stereo, int coarse, int32_t* matches12,
                                   float* points12);

/* Frame::UndistortKeyPoints (Frame.cc:730-763): cv::undistortPoints(pts, mK, mDistCoef, I, mK) on
 * the keypoint positions — OpenCV's five fixed-point iterations in double precision, radial (k1, k2, k3) and
 * tangential (p1, p2) terms.  K4 = fx, fy, cx, cy; dist = k1, k2, p1, p2[, k3] (ndist 4 or 5).  dist[0] == 0
 * copies the keypoints unchanged (:732-736).  kps_un may alias kps.  Returns 0 or a negative status. */
int orbfe_undistort_keypoints(orbfe_matcher* m, const orbfe_keypoint* kps, int n, const float* K4, const float* dist, int ndist,
                              orbfe_keypoint* kps_un);

/* MapPoint::ComputeDistinctiveDescriptors (MapPoint.cc:349-423) for `npoints` map points at
 * once: point p's observed descriptors are rows offsets[p] .. offsets[p+1] of desc (in the order the reference
 * collects them, MapPoint.cc:368-383).  best[p] = row (relative to offsets[p]) with the least median Hamming
 * distance to all of the point's rows (median = sorted[0.5 * (n - 1)], first row wins ties); -1 for a point
 * without observations.  Returns 0 or a negative status. */
int orbfe_distinctive_descriptors(orbfe_matcher* m, const uint8_t* desc, const int32_t* offsets, int npoints, int32_t* best);

/* ---- projection-guided searches -----------------------------------------------------
 * The Frame being searched: what Frame::GetFeaturesInArea (Frame.cc:640-706) and the SearchByProjection loops
 * read of it. */
typedef struct orbfe_grid_frame {
    const orbfe_keypoint* kps;     /* mvKeysUn (nleft == -1) or mvKeys followed by mvKeysRight */
    const uint8_t* desc;           /* mDescriptors (n x 32; right-camera rows after the left ones) */
    int n, nleft;                  /* Frame::N, Frame::Nleft */
    const uint8_t* blocked;        /* [n] mvpMapPoints[i] && mvpMapPoints[i]->Observations() > 0; NULL = none */
    const float* u_right;          /* mvuRight [n]; NULL = all -1 */
    const int32_t* left_to_right;  /* mvLeftToRightMatch [nleft]; NULL = all -1 */
    const int32_t* right_to_left;  /* mvRightToLeftMatch [n - nleft]; NULL = all -1 */
    float min_x, min_y;            /* Frame::mnMinX, mnMinY */
    float grid_w_inv, grid_h_inv;  /* mfGridElementWidthInv, mfGridElementHeightInv (64 x 48 grid, Frame.h:45-46) */
    const float* inv_level_sigma2; /* mvInvLevelSigma2 [n_levels]; only read by windows with chi2 != 0; NULL = none */
    int n_levels;
} orbfe_grid_frame;

/* One search window = one GetFeaturesInArea(u, v, r, min_level, max_level, right) call of the reference plus
 * what is matched inside it.  Windows are listed in the reference's loop order (a MapPoint seen by both cameras
 * of a fisheye rig gives two consecutive windows).  64 bytes. */
typedef struct orbfe_proj_query {
    float u, v, r;
    float ur;                      /* predicted right-image x; with check_ur: skip features with mvuRight > 0 and
                                      |ur - mvuRight| > r (ORBmatcher.cc:96-102, :2252-2258; only when nleft == -1) */
    float angle;                   /* angle of the source keypoint (rotation histogram) */
    int16_t min_level, max_level;  /* as passed to GetFeaturesInArea (max_level -1 = open) */
    uint8_t right;                 /* search the right-camera features */
    uint8_t obs;                   /* the MapPoint has Observations() > 0: once assigned, later windows skip that feature */
    uint8_t check_ur;
    uint8_t valid;                 /* 0 = the reference `continue`s before searching (not in view, bad, behind the camera ...) */
    uint8_t skip_if_prev_ratio_rejected; /* :135: a failed ratio test also skips the same MapPoint's right-camera search */
    uint8_t chi2;                  /* 1 = Fuse's reprojection gate (ORBmatcher.cc:1744-1770): skip a feature when
                                      ((u-x)^2 + (v-y)^2 [+ (ur-mvuRight)^2 if mvuRight >= 0]) * inv_level_sigma2[octave]
                                      exceeds 5.99 [7.8] */
    uint8_t skip_if_prev_empty;    /* skip this window when the PREVIOUS window's GetFeaturesInArea returned nothing: in
                                      SearchByProjection(Current, Last, ...) that `continue` (:2236-2237) also skips the same
                                      MapPoint's right-camera window */
    uint8_t pad;
    uint8_t desc[32];              /* pMP->GetDescriptor() */
} orbfe_proj_query;

/* The loop shared by SearchByProjection(Frame&, const vector<MapPoint*>&, th, ...) (ORBmatcher.cc:44-236: mode 1,
 * th = TH_HIGH, link_stereo = 1, rot_hist = 0) and SearchByProjection(Frame& Current, const Frame& Last, th, bMono)
 * (:2173-2389) / (Frame&, KeyFrame*, sAlreadyFound, th, ORBdist) (:2391-2510) (mode 0, th = TH_HIGH / ORBdist,
 * rot_hist = 1 (applied iff the matcher checks orientation)).  The KeyFrame-side searches are the same loop with
 * windows that never claim (obs = 0) or always claim: SearchByProjection(KeyFrame*, Scw, vpPoints, vpMatched, th,
 * ratioHamming) :473-585 / :588 (mode 0, th = floor(TH_LOW * ratioHamming), blocked = vpMatched non-NULL, obs = 1),
 * Fuse :1629-1821 (chi2 = 1) / :1823-1945 and both directions of SearchBySim3 :1947-2171 (obs = 0, read `choice`).
 *   mode 0: keep the nearest descriptor, accept dist <= th.
 *   mode 1: best and second best; accept dist <= th unless both sit on the same pyramid level and
 *           best > mfNNratio * second (:133-139).
 * assign[F->n]: index of the window whose MapPoint ends up in mvpMapPoints[i]; -1 = entry untouched;
 * -2 = set to NULL by the rotation check.  choice[nq] (nullable): feature chosen by each window or -1.
 * The windows are processed as if one after another (later windows see earlier assignments).
 * Returns nmatches as the reference counts it, or a negative status. */
int orbfe_search_by_projection(orbfe_matcher* m, const orbfe_grid_frame* F, const orbfe_proj_query* queries, int nq, int mode,
                               int th, int link_stereo, int rot_hist, int32_t* assign, int32_t* choice);

/* ORBmatcher::SearchForInitialization(Frame& F1, Frame& F2, vector<cv::Point2f>& vbPrevMatched, vector<int>& vnMatches12,
 * int windowSize) (ORBmatcher.cc:706-821), the matcher of Tracking::MonocularInitialization (Tracking.cc:2569).
 * kps1 / desc1 = F1.mvKeysUn / F1.mDescriptors [n1]; F2 = the frame whose grid is searched (kps = mvKeysUn, nleft = -1;
 * blocked, u_right and the stereo links are not read).  prev_matched [n1][2] = vbPrevMatched: read as the window
 * centres, and entries of matched features are overwritten with the matched F2 keypoint (:815-817).
 * matches12 [n1] = vnMatches12 (-1 = none).  Uses the matcher's mfNNratio and mbCheckOrientation.  Only level-0
 * keypoints of both frames take part; F1 features are processed as if one after another: a later feature with a
 * strictly smaller distance takes over an F2 feature, one with a larger or equal distance does not see it.
 * Returns nmatches or a negative status. */
int orbfe_search_for_initialization(orbfe_matcher* m, const orbfe_keypoint* kps1, const uint8_t* desc1, int n1,
                                    const orbfe_grid_frame* F2, float* prev_matched, int window_size, int32_t* matches12);

/* ---- BoW transform ------------------------------------------------------------------
 * DBoW2::TemplatedVocabulary<FORB::TDescriptor, FORB> (= ORB_SLAM3::ORBVocabulary) flattened: node 0 is the
 * root, nodes 1.. in the order of the vocabulary text file (saveToTextFile / loadFromTextFile,
 * Thirdparty/DBoW2/DBoW2/TemplatedVocabulary.h:1350-1454): parent[i] < i, is_leaf[i], a 32-byte descriptor
 * and a weight per node.  Word ids are assigned to the leaves in that order (:1413-1420); children keep file
 * order, which decides distance ties (:1256). */
typedef struct orbfe_vocabulary orbfe_vocabulary;
enum { ORBFE_BOW_TF_IDF = 0, ORBFE_BOW_TF = 1, ORBFE_BOW_IDF = 2, ORBFE_BOW_BINARY = 3 };   /* DBoW2::WeightingType */
enum { ORBFE_BOW_NORM_NONE = 0, ORBFE_BOW_NORM_L1 = 1, ORBFE_BOW_NORM_L2 = 2 };             /* mustNormalize(): L1 for ORBvoc */

int orbfe_vocabulary_create(int k, int L, int weighting, int norm, int nnodes, const int32_t* parent, const uint8_t* is_leaf,
                            const uint8_t* descriptors, const double* weights, int device, orbfe_vocabulary** out);
/* The ORBvoc.txt format ("k L scoring weighting" + one "parent isLeaf d0..d31 weight" line per node). */
int orbfe_vocabulary_load_text(const char* path, int norm, int device, orbfe_vocabulary** out);
void orbfe_vocabulary_destroy(orbfe_vocabulary* v);
int orbfe_vocabulary_words(const orbfe_vocabulary* v);
int orbfe_vocabulary_nodes(const orbfe_vocabulary* v);
void* orbfe_vocabulary_stream(orbfe_vocabulary* v);          /* cudaStream_t of the vocabulary's launches */
int orbfe_vocabulary_synchronize(orbfe_vocabulary* v);
long long orbfe_vocabulary_launch_count(const orbfe_vocabulary* v);

/* TemplatedVocabulary::transform(features, BowVector&, FeatureVector&, levelsup) (:1137-1206), what
 * Frame::ComputeBoW (Frame.cc:721-728) / KeyFrame::ComputeBoW (KeyFrame.cc:105-114) call with levelsup = 4.
 * Outputs are the two std::map's in key order: words[*nwords] / values (BowVector, double, bit-identical to
 * the reference's sequential sums and normalisation) and node_ids[*nnodes] + CSR offsets[*nnodes + 1] +
 * indices (FeatureVector = orbfe_feature_vector).  Capacities: n (offsets: n + 1).  Features whose word has
 * weight <= 0 (stop words) appear in neither.  A leaf above level L - levelsup reports its own node id (the
 * reference leaves nid uninitialised there). */
int orbfe_bow_transform(orbfe_vocabulary* v, const uint8_t* desc, int n, int levelsup, uint32_t* words, double* values,
                        int* nwords, uint32_t* node_ids, int32_t* offsets, int32_t* indices, int* nnodes);
/* Same for `batch` frames whose descriptors are already in device memory (e.g. orbfe_device_results):
 * frame b has d_n[b] rows at d_desc + b * frame_stride.  Device outputs, per frame at strides n_max
 * (d_words, d_values, d_node_ids, d_indices), n_max + 1 (d_offsets) and 1 (d_nwords, d_nnodes).
 * Enqueued on orbfe_vocabulary_stream; the caller orders it after the producer of d_desc. */
int orbfe_bow_transform_batch_device(orbfe_vocabulary* v, const uint8_t* d_desc, ptrdiff_t frame_stride, const int* d_n, int n_max,
                                     int batch, int levelsup, uint32_t* d_words, double* d_values, int* d_nwords,
                                     uint32_t* d_node_ids, int32_t* d_offsets, int32_t* d_indices, int* d_nnodes);

/* Measured integer-pipe ceiling of the Hamming matchers: sustained 32-bit POPC instructions per
 * second of device `device` (register-only micro-benchmark, ~10 ms).  One 256-bit descriptor pair
 * costs 8 POPC, so popc_per_s / 8 is the roofline of the matching kernels in pairs/s. */
int orbfe_measure_popc_peak(int device, double* popc_per_s);

const char* orbfe_strerror(int status);
const char* orbfe_last_cuda_error(void);
/* "orbfe <version> sm_100a" */
const char* orbfe_version(void);

#ifdef __cplusplus
}
#endif
#endif /* ORBFE_H_ */
