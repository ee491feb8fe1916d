/* include/orbx.h -- C ABI of the B200-native ORB front-end (liborbx_b200.so).
 *
 * This is the drop-in boundary for ONE hot path of ORB-SLAM3 (reference: zhoujoey/ORB-SLAM3-Mac,
 * cites are file:line relative to the reference tree):
 *
 *   orbx_*   replaces  ORB_SLAM3::ORBextractor            include/ORBextractor.h:43-109
 *                      ORBextractor::operator()           src/ORBextractor.cc:1068-1150
 *            called from Frame::ExtractORB                src/Frame.cc:410-417
 *   orbm_*   replaces  the ORBmatcher::DescriptorDistance loops (src/ORBmatcher.cc:2353-2369)
 *            inside SearchByProjection / SearchForInitialization (src/ORBmatcher.cc:48-218,
 *            710-825, 1965-2181) and the stereo matchers in src/Frame.cc:802-885, 1128-1153
 *            (declared in include/orbm.h).
 *
 * Plain C, POD only: no C++ exceptions, no OpenCV or torch types cross this boundary.
 * All work runs in hand-written sm_100a CUDA kernels; there is NO CPU fallback -- if no CUDA
 * device is usable, orbx_create fails with ORBX_E_CUDA.
 *
 * Threading: one handle == private device buffers + CUDA streams.  A handle is not re-entrant;
 * distinct handles are fully concurrent (the reference runs the left/right extractors on two
 * threads, src/Frame.cc:109-112).
 */
#ifndef ORBX_H
#define ORBX_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define ORBX_ABI_VERSION 1

typedef enum orbx_status {
    ORBX_OK = 0,
    ORBX_E_EMPTY_IMAGE = -1, /* reference: `return -1` on an empty image, ORBextractor.cc:1072-1073 */
    ORBX_E_BAD_ARG = -2,     /* NULL pointers, non-positive sizes, nlevels out of range, ... */
    ORBX_E_BAD_TYPE = -3,    /* reference: assert(image.type()==CV_8UC1), ORBextractor.cc:1076 */
    ORBX_E_CAPACITY = -4,    /* caller's output capacity < number of keypoints; nothing truncated */
    ORBX_E_TOO_LARGE = -5,   /* image/batch exceeds what the handle was created for */
    ORBX_E_UNSUPPORTED = -6, /* shape the reference itself cannot process (nIni==0 portrait
                                levels: division by zero at ORBextractor.cc:541-543; levels with
                                fewer than one 30-px FAST cell) or beyond the documented limits */
    ORBX_E_CUDA = -7,        /* CUDA runtime failure; see orbx_last_error_string() */
    ORBX_E_BUSY = -8         /* async slot still in flight */
} orbx_status;

/* Same field order and size as cv::KeyPoint (28 bytes), so a C++ wrapper can memcpy. */
typedef struct orbx_keypoint {
    float x, y;       /* pt, in level-0 pixel coordinates (scaled by mvScaleFactor[octave]) */
    float size;       /* (int)(31 * mvScaleFactor[octave]) */
    float angle;      /* degrees in [0,360), cv::fastAtan2 polynomial */
    float response;   /* FAST score */
    int32_t octave;   /* pyramid level */
    int32_t class_id; /* -1 */
} orbx_keypoint;

/* The five ORBextractor constructor arguments (ORBextractor.cc:408-409; read from yaml at
 * src/Tracking.cc:200-204) plus sizing of the device-resident buffers. */
typedef struct orbx_config {
    int32_t nfeatures;
    float scale_factor;
    int32_t nlevels;     /* 1..ORBX_MAX_LEVELS */
    int32_t ini_th_fast;
    int32_t min_th_fast;
    int32_t max_width;   /* largest image the handle must accept */
    int32_t max_height;
    int32_t max_batch;   /* frames per launch of the batched entry points (>=1) */
    int32_t n_slots;     /* pipeline depth of the async entry points (>=1; 0 -> 2) */
    int32_t device;      /* CUDA device ordinal */
    uint32_t flags;      /* ORBX_FLAG_* */
} orbx_config;

#define ORBX_MAX_LEVELS 16
#define ORBX_FLAG_NONE 0u
#define ORBX_FLAG_STAGE_TIMING 1u /* record CUDA events at the stage boundaries (bench.py) */
#define ORBX_FLAG_NO_GRAPH 2u     /* never replay a CUDA graph for small batches (A/B measurements) */
#define ORBX_FLAG_NO_FORK 4u      /* capture that graph as one chain instead of one branch per pyramid level */
#define ORBX_GRAPH_MAX_FRAMES 4   /* batches up to this size replay a captured graph of the kernels */

typedef struct orbx_handle orbx_handle;

/* ---- lifetime ------------------------------------------------------------------------ */
int orbx_create(const orbx_config* cfg, orbx_handle** out);
int orbx_destroy(orbx_handle* h);
const char* orbx_last_error_string(void); /* thread-local, never NULL */
int orbx_abi_version(void);

/* Per-level tables: the getters of include/ORBextractor.h:61-81 (GetScaleFactors,
 * GetInverseScaleFactors, GetScaleSigmaSquares, GetInverseScaleSigmaSquares) and
 * mnFeaturesPerLevel.  Each array has nlevels entries; any pointer may be NULL. */
int orbx_get_tables(const orbx_handle* h, float* scale, float* inv_scale, float* sigma2,
                    float* inv_sigma2, int32_t* n_per_level);

/* Output capacity (keypoints per frame) that can never overflow for this handle:
 * nfeatures + nlevels * max(3, 4*nIni) rounded up (SURVEY.md 8 E3 bound). */
int orbx_max_keypoints(const orbx_handle* h);

/* ---- extraction: host buffers in, host buffers out (the reference-facing call) -------- */
/* ORBextractor::operator()(image, mask, keypoints, descriptors, vLappingArea).
 * img: h rows of w bytes, row stride stride_bytes, 8-bit single channel.
 * lap_x0/lap_x1 = vLappingArea[0..1]: keypoints with lap_x0 <= pt.x <= lap_x1 are written from
 * the BACK of the arrays in reverse order, the others from the front (ORBextractor.cc:1135-1144).
 * kps/desc: caller-owned, cap records / cap*32 bytes.  *n_out = number of keypoints,
 * *mono_index = the reference's return value.  Synchronous. */
int orbx_extract(orbx_handle* h, const uint8_t* img, int w, int hgt, size_t stride_bytes,
                 int lap_x0, int lap_x1, orbx_keypoint* kps, uint8_t* desc, int cap, int* n_out,
                 int* mono_index);

/* Batch of nframes (<= max_batch) images of identical size; imgs[i] as above.
 * lap: NULL (-> {0,0} for every frame) or 2*nframes ints.  Outputs are strided by cap:
 * frame i owns kps[i*cap .. i*cap+n_out[i]) and desc[i*cap*32 ..).  status[i] (may be NULL)
 * receives a per-frame orbx_status.  Synchronous. */
int orbx_extract_batch(orbx_handle* h, const uint8_t* const* imgs, int nframes, int w, int hgt,
                       size_t stride_bytes, const int* lap, orbx_keypoint* kps, uint8_t* desc,
                       int cap, int* n_out, int* mono_index, int* status);

/* Pipelined form: submit returns after enqueueing H2D + kernels + D2H on the slot's stream;
 * wait blocks until that slot's outputs are complete in the caller's buffers.  Host buffers
 * must stay valid (and should be pinned, see orbx_host_alloc) until wait returns. */
int orbx_submit_batch(orbx_handle* h, int slot, const uint8_t* const* imgs, int nframes, int w,
                      int hgt, size_t stride_bytes, const int* lap, orbx_keypoint* kps,
                      uint8_t* desc, int cap, int* n_out, int* mono_index, int* status);
int orbx_wait(orbx_handle* h, int slot);

/* Pinned host memory for the pipelined form. */
void* orbx_host_alloc(size_t bytes);
void orbx_host_free(void* p);

/* ---- extraction on device-resident frames (kernel-only timing; multi-GPU result gather) -- */
/* d_imgs: device pointer to nframes contiguous images (row pitch pitch_bytes, frame stride
 * frame_stride_bytes).  Results stay on the device in the slot's result slab:
 *   d_header  int32[nframes][4] = {n, mono_index, status, 0}
 *   d_kps     orbx_keypoint[nframes][orbx_max_keypoints()]
 *   d_desc    uint8[nframes][orbx_max_keypoints()][32]
 * Asynchronous on the slot's stream; orbx_wait(h, slot) synchronises. */
int orbx_extract_device(orbx_handle* h, int slot, const uint8_t* d_imgs, int nframes, int w,
                        int hgt, size_t pitch_bytes, size_t frame_stride_bytes, const int* lap,
                        const int32_t** d_header, const orbx_keypoint** d_kps,
                        const uint8_t** d_desc);
/* The CUDA stream (cudaStream_t) of a slot, for event timing by the caller. */
void* orbx_slot_stream(orbx_handle* h, int slot);

/* ---- multi-GPU: the result gather fused into the producing kernel -------------------
 * Frames are sharded over GPUs as independent units; the only exchange is gathering the results.  A result
 * sink is a slab on the gathering GPU -- [frames][4] headers {n, mono_index, status, 0}, [frames][cap]
 * keypoints, [frames][cap][32] descriptors, cap = orbx_max_keypoints() -- that OTHER GPUs' k_describe
 * kernels write into directly: every result row goes to the local slab and, in the same kernel, to frame
 * first_frame + f of the sink (peer memory over NVLink when the sink lives on another GPU).  No collective
 * follows the kernel; the gathering side reads the slab after the producers' streams have completed (stream
 * synchronisation + a host barrier, or any event the application already has).
 *   same process:   orbx_sink_create on the gathering device, orbx_set_result_sink on every handle (peer access
 *                   is enabled on demand);
 *   one process per GPU: orbx_sink_create + orbx_sink_export on the gathering rank, the 64-byte CUDA IPC handle
 *                   travels through the application's own channel, orbx_sink_open on the other ranks. */
typedef struct orbx_sink orbx_sink;
int orbx_sink_create(int device, int frames, int cap, orbx_sink** out);
int orbx_sink_export(orbx_sink* s, uint8_t handle[64]);
int orbx_sink_open(int device, const uint8_t handle[64], int frames, int cap, orbx_sink** out);
int orbx_sink_destroy(orbx_sink* s);
/* Device pointers of the three arrays (valid on the creating / opening device). */
int orbx_sink_pointers(orbx_sink* s, void** header, void** kps, void** desc);
/* Host copy of one gathered frame (synchronises the sink's device first). */
int orbx_sink_read_frame(orbx_sink* s, int frame, int32_t header[4], orbx_keypoint* kps, uint8_t* desc, int rows);
/* Runs of `slot` write frame f of a batch to sink frame first_frame + f from now on (sink NULL: off).
 * The sink must hold first_frame + max_batch frames of the handle's capacity. */
int orbx_set_result_sink(orbx_handle* h, int slot, orbx_sink* sink, int first_frame);

/* Completion signal and back-pressure of the fused gather.  The region [first_frame, first_frame + max_batch) that a
 * slot writes owns two words of the sink: `flag` and `ack`.  After batch e of the slot (e = 1, 2, ... since the sink
 * was attached) is completely in the sink -- rows, headers -- the producer publishes e at flag (one tiny kernel behind
 * k_describe on the slot's stream, release at system scope).  A consumer on the gathering GPU
 *     orbx_sink_wait(sink, first_frame, e, stream, timeout_ms)
 * either blocks until batch e has arrived (stream == NULL; ORBX_E_BUSY after timeout_ms) or orders everything enqueued
 * later on `stream` (a cudaStream_t) behind its arrival, reads the region in frame order, and then
 *     orbx_sink_release(sink, first_frame, e, stream)
 * With ORBX_SINK_BACKPRESSURE the producer's k_describe of batch e + 1 does not start before batch e was released, so a
 * slow consumer is never overwritten (a producer that waits for more than 2 s gives up and sets the slot's timed_out
 * flag, see orbx_sink_epoch).  While a producer is parked, the consumer's side must not do anything that serialises
 * behind running kernels: first launches of not-yet-loaded kernels (lazy module loading; the library pre-loads its
 * own), cudaMalloc / cudaFree, pageable-memory copies.  Without the flag the producer never waits (the round-1 behaviour: the application
 * orders consumer and producers itself). */
#define ORBX_SINK_BACKPRESSURE 1
int orbx_set_result_sink_ex(orbx_handle* h, int slot, orbx_sink* sink, int first_frame, int flags);
int orbx_sink_wait(orbx_sink* s, int first_frame, uint32_t epoch, void* stream, int timeout_ms);
int orbx_sink_release(orbx_sink* s, int first_frame, uint32_t epoch, void* stream);
/* Batches the slot has sent to its sink so far (= the epoch of the last one); timed_out (may be NULL) = a back-pressure
 * wait of this slot has given up since the sink was attached. */
int orbx_sink_epoch(orbx_handle* h, int slot, uint32_t* epoch, int* timed_out);

/* ---- pyramid access: ORBextractor::mvImagePyramid (include/ORBextractor.h:83), read by
 * Frame::ComputeStereoMatches (src/Frame.cc:809,899-918) ------------------------------------- */
int orbx_level_size(const orbx_handle* h, int w, int hgt, int level, int* lw, int* lh);
/* Copies level `level` of frame `frame` of the last batch run on `slot` to host memory.
 * pad = 0 copies the level itself; pad = 19 reproduces the reference's padded buffer
 * (EDGE_THRESHOLD border, BORDER_REFLECT_101) of (lw+38) x (lh+38). */
int orbx_get_pyramid_level(orbx_handle* h, int slot, int frame, int level, int pad, uint8_t* dst,
                           size_t dst_stride);

/* Device-resident view of one frame's pyramid of the last batch run on `slot`: what a consumer on
 * the same GPU (orbm_compute_stereo_matches: the SAD refinement of src/Frame.cc:888-963 reads
 * BOTH extractors' mvImagePyramid[octave]) uses instead of copying levels to the host.  The
 * pointers stay valid until the next run on that slot; no border is materialised (readers apply
 * BORDER_REFLECT_101 themselves, as orbx_get_pyramid_level(pad) does). */
typedef struct orbx_pyramid_view {
    const uint8_t* data[ORBX_MAX_LEVELS];  /* device pointers */
    int32_t pitch[ORBX_MAX_LEVELS];        /* bytes per row */
    int32_t width[ORBX_MAX_LEVELS];
    int32_t height[ORBX_MAX_LEVELS];
    int32_t nlevels;
    int32_t device;
} orbx_pyramid_view;
int orbx_device_pyramid(orbx_handle* h, int slot, int frame, orbx_pyramid_view* out);

/* ---- stage taps for the parity tests (results of the last batch run on `slot`) ------------ */
/* FAST candidates of one level in the reference's order (cells row-major, raster inside a
 * cell): int32 triplets {x, y, response}, coordinates relative to (16,16) as
 * vToDistributeKeys in ORBextractor.cc:776-854.  Returns the count (or <0). */
int orbx_debug_candidates(orbx_handle* h, int slot, int frame, int level, int32_t* xyr, int cap);
/* Parity tap of the descriptor stage's cos / sin (computeOrbDescriptor's (float)cos(angle), (float)sin(angle),
 * src/ORBextractor.cc:113-114): the kernel's own double-precision evaluation against CUDA's sincos() on EVERY float in
 * [lo, hi] (0 <= lo <= hi); *n_bad = values whose float cos or sin differs, *first_bad = the smallest of them. */
int orbx_debug_sincos_check(int device, float lo, float hi, unsigned long long* n_bad, float* first_bad);
/* Quadtree survivors of one level in DistributeOctTree's output order (ORBextractor.cc:738-758):
 * int32 triplets {x, y, response}, same coordinate convention.  Returns the count (or <0). */
int orbx_debug_survivors(orbx_handle* h, int slot, int frame, int level, int32_t* xyr, int cap);

/* ---- input stage: what the reference does to an image before ExtractORB ----------
 * cvtColor(mImGray, mImGray, CV_{RGB,BGR,RGBA,BGRA}2GRAY)   src/Tracking.cc:354-379, 425-438, 479-492
 * cv::remap(im, imRect, M1, M2, cv::INTER_LINEAR)            Examples/Stereo/stereo_euroc.cc:170-171
 * fused in front of pyramid level 0 on the device.  Both settings persist on the handle and apply to
 * every later extract / submit / extract_device call, whose `img` is then the RAW image (stride in
 * bytes, >= w * channels); level 0 of the pyramid is the rectified gray image. */
#define ORBX_IN_GRAY8 0 /* CV_8UC1 (default) */
#define ORBX_IN_BGR8 1  /* CV_8UC3, mbRGB == false */
#define ORBX_IN_RGB8 2  /* CV_8UC3, mbRGB == true */
#define ORBX_IN_BGRA8 3 /* CV_8UC4 */
#define ORBX_IN_RGBA8 4
/* gray_bits: 15 = the luma coefficients of OpenCV >= 4 (9798, 19235, 3735, >> 15), 14 = those of
 * OpenCV 2.4/3.x (4899, 9617, 1868, >> 14). */
int orbx_set_input_format(orbx_handle* h, int format, int gray_bits);
/* map_x / map_y: the CV_32FC1 maps of cv::initUndistortRectifyMap (stereo_euroc.cc:117-118), w x hgt
 * floats each, rows map_stride_floats apart, for images of exactly that size; converted once to
 * OpenCV's fixed point (cvRound(map * 32)) and kept on the device.  NULL maps switch the stage off.
 * Source pixels outside the image read 0 (cv::remap's default BORDER_CONSTANT). */
int orbx_set_rectify_maps(orbx_handle* h, const float* map_x, const float* map_y, int w, int hgt,
                          size_t map_stride_floats);

/* With ORBX_FLAG_STAGE_TIMING: device milliseconds of the last run on `slot`, measured with CUDA
 * events on the slot's stream: ms[0..4] = pyramid, FAST, quadtree, assemble, describe; ms[5] =
 * all kernels.  Returns the number of values written. */
int orbx_last_stage_ms(orbx_handle* h, int slot, float* ms, int n);

/* How many kernels the last submit on `slot` launched (bench.py's gpu_launches). */
int orbx_last_launch_count(const orbx_handle* h, int slot);

#ifdef __cplusplus
}
#endif
#endif /* ORBX_H */
