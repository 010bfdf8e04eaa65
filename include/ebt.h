/* ebt.h — C-ABI of the B200-native event-based feature tracker.
 *
 * Drop-in boundary for the reference's tracker API
 * (/root/reference/implementation/feature_tracker/include/feature_tracker/
 * feature_detector.h:33-120 and patch.h:14-130).  The reference has no FFI of
 * its own; each entry point below names the C++ member it replaces.  The C++
 * facade in include/ebt_tracker.hpp rebuilds tracker::FeatureDetector /
 * tracker::Patch on top of these calls.
 *
 * Conventions: plain pointers and sizes, caller owns every buffer passed in
 * (copied during the call), outputs are copied into caller arrays, no device
 * pointer escapes except through the *_device entry points.  Every function
 * returns 0 on success or a negative EBT_E_* code; ebt_last_error() gives the
 * message.  There is no CPU fallback: without a CUDA device ebt_create fails.
 * A context is bound to one CUDA device and one stream and is not thread-safe;
 * different contexts may be driven from different host threads.
 */
#ifndef EBT_H_
#define EBT_H_

#include <stddef.h>
#include <stdint.h>

#include "ebt_types.h"

#ifdef __cplusplus
extern "C" {
#endif

#define EBT_OK 0
#define EBT_E_INVALID -1  /* bad argument */
#define EBT_E_CUDA -2     /* CUDA runtime error (message holds cudaGetErrorString) */
#define EBT_E_CAPACITY -3 /* a fixed capacity was exceeded */
#define EBT_E_NODEVICE -4 /* no usable CUDA device */

typedef struct ebt_ctx ebt_ctx;

/* tracker::DetectorParams (feature_detector.h:10-31) + tracker::OptimizerParams
 * (optimizer.h:16-27) + the constants hard-coded in tracker::Patch
 * (patch.h:113-114, patch.cpp:22,298), as far as the hot path reads them. */
typedef struct ebt_params
{
	int32_t width, height;          /* DetectorParams::imageSize */
	int32_t max_patches;            /* DetectorParams::maxPatches (100) */
	int32_t init_num_events;        /* DetectorParams::initNumEvents (75) */
	int32_t min_num_events;         /* Patch::minNumOfEvents_ (100) */
	int32_t max_num_events;         /* Patch::maxNumOfEvents_ (300) */
	int32_t batch_stride;           /* Patch::isReady literal (30) */
	int32_t max_iterations;         /* OptimizerParams::maxNumIterations (10) */
	int32_t border_margin;          /* updateNumOfEvents literal (5) */
	int32_t record_capacity;        /* optimisation records kept per push (device) */
	double association_distance;    /* DetectorParams::associationDistance (5) */
	double huber_loss;              /* OptimizerParams::huberLoss (0.3) */
	double optimizer_threshold;     /* OptimizerParams::optimizerThreshold (0.6) */
} ebt_params;

/* Fills *p with the reference defaults for a width x height sensor. */
int ebt_default_params(ebt_params* p, int width, int height);

/* FeatureDetector::FeatureDetector(const DetectorParams&) (feature_detector.cpp:9-13). */
int ebt_create(const ebt_params* params, int device, ebt_ctx** out);
int ebt_destroy(ebt_ctx* ctx);
const char* ebt_last_error(const ebt_ctx* ctx); /* ctx may be NULL: last create error */

/* FeatureDetector::newImage (feature_detector.cpp:493-541), split at the two
 * places where the reference calls OpenCV's detector and LK tracker, whose
 * results the host injects:
 *   begin:  extractPatches (:543-566) with corners = detectFeatures(image)
 *           (cornersXY: n x (x, y) doubles) -> log image, Sobel gradients,
 *           new template grid, associatePatches (:630-664);
 *   get_uninit / set_flow: FlowEstimator::getFlowPatches (flow_estimator.cpp:
 *           27-82): query the patches LK would run on, hand back nextPoint;
 *   end:    updateNumOfEvents for all (:499-502), archive lost patches,
 *           release unused template grids (:512-535).
 * gray: height*width uint8, row-major (common::ImageSample::value, CV_8U). */
int ebt_new_frame_begin(ebt_ctx* ctx, const uint8_t* gray, int64_t ts_us,
						const double* cornersXY, int n_corners);
/* Same as ebt_new_frame_begin with the gray frame already on ctx's device. */
int ebt_new_frame_begin_device(ebt_ctx* ctx, const uint8_t* d_gray, int64_t ts_us,
							   const double* cornersXY, int n_corners);
int ebt_get_uninit(ebt_ctx* ctx, int32_t* track_ids, double* cornersXY, int cap, int* n_out);
int ebt_set_flow(ebt_ctx* ctx, int32_t track_id, int found, float next_x, float next_y);
/* ebt_set_flow for n patches in one call; nextXY = n (x, y) float pairs. */
int ebt_set_flows(ebt_ctx* ctx, const int32_t* track_ids, const int32_t* found, const float* nextXY, int n);
int ebt_new_frame_end(ebt_ctx* ctx);
/* ebt_new_frame_end for the detectors of n_streams streams (configs[3]) from one host
 * thread: the window-size kernels and state copies of all of them are queued first,
 * each on its context's CUDA stream, then waited for.  Errors are reported on ctxs[0]. */
int ebt_new_frame_end_multi(ebt_ctx* const* ctxs, int n_streams);

/* ---- per-frame re-arm on the device ------------------------
 * What the reference delegates to OpenCV inside newImage, as sm_100a kernels
 * (csrc/ebt_frontend.cu); the algorithm is OpenCV 4.x's, pinned by cv2 golden
 * vectors (tests/golden/frontend_cv2.npz).
 *
 * DetectorParams::qualityLevel / minDistance (feature_detector.h:12-13) and
 * FlowEstimatorParams::numLevels (flow_estimator.h:13); blockSize 3, Harris
 * k 0.04, patchExtent 12 and the 21x21 LK window are fixed as in the
 * reference's defaults. */
typedef struct ebt_frontend_params
{
	double quality_level; /* 0.01 */
	double min_distance;  /* 10 */
	int32_t lk_levels;    /* maxLevel 3 */
	int32_t lk_max_count; /* TermCriteria default of calcOpticalFlowPyrLK: 30 */
	double lk_epsilon;    /* 0.01 */
} ebt_frontend_params;
int ebt_default_frontend_params(ebt_frontend_params* p);
int ebt_set_frontend_params(ebt_ctx* ctx, const ebt_frontend_params* p);

/* FeatureDetector::detectFeatures (feature_detector.cpp:568-583) =
 * cv::goodFeaturesToTrack(gray, W*H/625, qualityLevel, minDistance, mask_, 3,
 * true, 0.04) with the mask of reset() (:34-43).  cornersXY receives up to cap
 * (x, y) pairs in OpenCV's order (strongest first); *n_out = number found. */
int ebt_detect_features(ebt_ctx* ctx, const uint8_t* gray, double* cornersXY, int cap, int* n_out);

/* FlowEstimator::getFlow (flow_estimator.cpp:84-103) for n points at once =
 * cv::calcOpticalFlowPyrLK(prev, cur, pts, next, status, err, {21, 21},
 * numLevels).  ptsXY / nextXY: n (x, y) float pairs; status: n bytes. */
int ebt_lk_flow(ebt_ctx* ctx, const uint8_t* prev_gray, const uint8_t* cur_gray, const float* ptsXY,
				int n, float* nextXY, uint8_t* status);

/* FeatureDetector::newImage (feature_detector.cpp:493-541) in one call:
 * detectFeatures -> extractPatches/associatePatches -> FlowEstimator::addImage
 * + getFlowPatches on the previous/current frame pair kept on the device ->
 * updateNumOfEvents -> archive.  Equivalent to ebt_new_frame_begin /
 * ebt_get_uninit / ebt_set_flows / ebt_new_frame_end with the corners and LK
 * results computed by the kernels above. */
int ebt_new_image(ebt_ctx* ctx, const uint8_t* gray, int64_t ts_us);

/* FeatureDetector::getFeatures (feature_detector.h:69): corners_ of the latest
 * frame, (x, y) pairs; *n_out = number available (may exceed cap). */
int ebt_get_features(ebt_ctx* ctx, double* cornersXY, int cap, int* n_out);

/* FeatureDetector::updatePatches (feature_detector.cpp:585-619) for n events in
 * stream order (the reference takes one event per call).  Host buffer; the
 * call copies it to the device, runs the stream kernel and returns when the
 * resulting track state is back on the host; the optimisation records of the
 * push are still travelling then (their copy overlaps the caller's next image)
 * and are complete whenever a getter that reads them returns.
 * *n_optimizations (may be NULL) receives the number of Optimizer::optimize
 * equivalents performed. */
int ebt_push_events(ebt_ctx* ctx, const ebt_event* events, int64_t n, int64_t* n_optimizations);

/* Same, events already resident on the context's device (no H2D copy). */
int ebt_push_events_device(ebt_ctx* ctx, const ebt_event* d_events, int64_t n,
						   int64_t* n_optimizations);

/* configs[3] (64 independent streams): the reference runs one FeatureDetector per
 * stream; this is FeatureDetector::updatePatches (feature_detector.cpp:585-619)
 * for every event of n_streams detectors' batches in ONE prefilter launch and
 * ONE stream-kernel launch, driven by one host thread with one wait for all of
 * them.  events[i] / n_events[i] is stream i's batch for ctxs[i] (host buffers,
 * or buffers on the contexts' device when device_resident != 0); results per
 * stream are exactly those of ebt_push_events on that context alone.  All
 * contexts must live on the same device and have equal image size and solver
 * parameters; errors are reported on ctxs[0].  n_optimizations (may be NULL)
 * receives n_streams counts. */
int ebt_push_events_multi(ebt_ctx* const* ctxs, int n_streams, const ebt_event* const* events,
						  const int64_t* n_events, int device_resident, int64_t* n_optimizations);

/* ---- one stream's features over several GPUs (north_star: "partitioned ... by
 * feature batch or stream with only a host-side gather of tracks") ------------
 * The reference has one FeatureDetector per stream on one core; there is no
 * counterpart.  Between two images a patch's evolution depends only on its own
 * state and the ordered event stream (feature_detector.cpp:587-617); the only
 * cross-feature steps are associatePatches and the archive at image time
 * (:512-535, 630-664).  So G contexts (one per GPU) are fed the SAME images and
 * the SAME event batches; context `index` of `count` tracks the patches whose
 * track id = index (mod count) and carries the others along as list entries only.
 * After every ebt_push_events the caller gathers the shards' patch states on the
 * host (ebt_get_patch_state: n x ebt_sizeof_patch_state() opaque bytes, list order)
 * and hands all of them to every shard (ebt_merge_patch_state), so that the next
 * image's association, window sizes and archive decisions are the same everywhere
 * — and the same as on one GPU.  Optimisation records, trajectories and patch
 * products of a track live on its owner shard.  No GPU-to-GPU traffic, no collective. */
int ebt_set_shard(ebt_ctx* ctx, int index, int count);
size_t ebt_sizeof_patch_state(void);
int ebt_get_patch_state(ebt_ctx* ctx, void* out, int cap, int* n_out);
int ebt_merge_patch_state(ebt_ctx* ctx, const void* const* states, int n_shards, int n_patches);

/* FeatureDetector::getPatches / getArchivedPatches (feature_detector.h:71-76):
 * snapshots in list order.  *n_out = number available (may exceed cap). */
int ebt_get_tracks(ebt_ctx* ctx, ebt_track* out, int cap, int archived, int* n_out);

/* Patch::getIntegratedNabla / getCompenatedIntegratedNabla / getPredictedNabla
 * (patch.h:49-50, 67): 625 doubles, row-major at<double>(y,x).
 * which: 0 = integratedNabla_, 1 = motionCompensatedIntegratedNabla_,
 * 2 = predictedNabla_ (1 and 2 need ebt_set_products). */
int ebt_get_patch(ebt_ctx* ctx, int32_t track_id, int which, double* out625);

/* The two Patch products that only feed getters: Patch::integrateMotion-
 * CompensatedEvents (patch.cpp:87-130, called by Optimizer::optimize,
 * optimizer.cpp:182) and Patch::warpImage (patch.cpp:132-154, called after
 * every optimisation, feature_detector.cpp:615, and for every patch at image
 * time when DetectorParams::drawImages is set, :504-510).
 * mode 0 (default): not computed; 1: after every optimisation (the reference
 * with drawImages = false); 2: also at image time (drawImages = true). */
int ebt_set_products(ebt_ctx* ctx, int mode);

/* OptimizerParams::drawCostMap (feature_tracker/optimizer.h:18) and
 * Optimizer::drawCostMap (ft/src/optimizer.cpp:32-60, called at :191-197): after
 * every optimisation the solve's functor is evaluated on the 11x11 grid of
 * integer translation offsets around the solved warp; the map holds the L2 norm
 * of the 625 residuals.  121 extra pixel passes per optimisation ("it is too
 * slow", optimizer.cpp:194): off by default, as in the reference.
 * ebt_get_cost_map replaces Patch::getCostMap (patch.h:56): out121 row-major,
 * map(y + 5, x + 5); zeros until the patch was optimised with the map on.
 * ebt_get_batch_cost_maps: the maps of the last ebt_optimize_batch(_device)
 * call made with the map on, n x 121. */
int ebt_set_cost_map(ebt_ctx* ctx, int on);
int ebt_get_cost_map(ebt_ctx* ctx, int32_t track_id, double* out121);
int ebt_get_batch_cost_maps(ebt_ctx* ctx, double* out, int64_t n);

/* Tracker -> visual odometry hand-off.  For every live patch,
 * in list order: the landmark visual_odometry::Keyframe::Keyframe stores
 * (vo/src/keyframe.cpp:5-14: trackId -> Patch::toCorner()) and, when cam is
 * not NULL, its bearing vector common::CameraModel::unproject(corner)
 * (common/include/common/camera_model.h:79-114: 10 fixed-point iterations of the
 * radial/tangential model, then normalisation), as the VO front end computes
 * it per landmark (vo/src/visual_odometry.cpp:234, 367-369, 516-518).  Computed
 * on the device from the resident patch state.  cornersXY: 2 doubles per patch,
 * bearingsXYZ: 3 doubles per patch (may be NULL); *n_out = patches available. */
int ebt_get_landmarks(ebt_ctx* ctx, const ebt_camera* cam, int32_t* track_ids, double* cornersXY,
					  double* bearingsXYZ, int cap, int* n_out);
/* CameraModel::unproject for n arbitrary pixel positions (same kernel). */
int ebt_unproject(ebt_ctx* ctx, const ebt_camera* cam, const double* pointsXY, int n, double* bearingsXYZ);

/* Patch::getTrajectory (patch.h:58): (x, y, t_us) triples. */
int ebt_get_trajectory(ebt_ctx* ctx, int32_t track_id, double* out_xyt, int cap, int* n_out);

/* Patch::getFinalCosts (patch.h:65). */
int ebt_get_final_costs(ebt_ctx* ctx, int32_t track_id, double* out, int cap, int* n_out);

/* Optimisation records since the last ebt_clear_records, in the order the
 * reference would have executed them (Optimizer::getFinalCosts,
 * optimizer.h:47, is the cost/ts projection of this list). */
int ebt_get_records(ebt_ctx* ctx, ebt_opt_record* out, int64_t cap, int64_t* n_out);
int ebt_clear_records(ebt_ctx* ctx);

/* Latest frame's gradients (FeatureDetector::gradX_/gradY_), width*height each. */
int ebt_get_gradients(ebt_ctx* ctx, double* gx, double* gy);

/* ---- event ingestion ----------------------------------------
 * tools::Davis240cReader::getEventSample (tools/dataset_reader/src/
 * davis240c_reader.cpp:60-92) for a whole text buffer on the device: lines
 * "t_sec x y p" -> ebt_event {t_us = (int64)(stod(t_sec) * 1e6), x, y,
 * polarity = p ? +1 : -1}.  final_chunk = 0: an unterminated last line is
 * left unparsed and *consumed (may be NULL) tells how many bytes were used, so
 * a file can be streamed in chunks (the reference reads 1e6 lines at a time,
 * :14); final_chunk = 1: the last line needs no newline (std::getline).
 * *n_out = lines found (EBT_E_CAPACITY if > cap).  A malformed line, a sign
 * other than 0/1 (the reference throws, :85-88) or a timestamp whose decimal
 * mantissa is >= 2^53 is EBT_E_INVALID naming the line. */
int ebt_parse_events_text(ebt_ctx* ctx, const char* text, size_t len, int final_chunk, ebt_event* out,
						  int64_t cap, int64_t* n_out, size_t* consumed);

/* ebt_parse_events_text for callers that must not allocate.  No counterpart in
 * the reference: added for code that
 * cannot call cudaMalloc (set-up under CUDA graph capture, tight serving loops,
 * harnesses that forbid it).  Same kernels and same results as
 * ebt_parse_events_text, but text and output are DEVICE buffers owned by the
 * caller (d_text 16-byte aligned, on `device`), and the parser's scratch (tile
 * line counts, totals, error word) is a fixed file-scope __device__ array of
 * 65 792 bytes that every GPU loading the library holds until exit.  The call
 * issues kernels, small memsets and two small D2H copies on `cuda_stream`
 * (a cudaStream_t, NULL = default stream) and waits for them; it calls no
 * cudaMalloc / cudaFree / cudaHostAlloc and no allocating library routine, and
 * needs no ebt_ctx.
 * Limit: len <= EBT_STATIC_PARSE_MAX_BYTES per call (EBT_E_CAPACITY beyond
 * it; stream longer files in chunks with final_chunk = 0 and *consumed).
 * The scratch is shared by all calls on a device and keeps the last call's
 * contents; the entry re-initialises what must start clean, but two calls that
 * overlap on the GPU (different host threads or streams) race on it.
 * errmsg (may be NULL) receives the message of a failure, errmsg_cap bytes. */
#define EBT_STATIC_PARSE_MAX_BYTES (64u << 20)
int ebt_parse_events_text_static(int device, void* cuda_stream, const char* d_text, size_t len, int final_chunk,
								 ebt_event* d_out, int64_t cap, int64_t* n_out, size_t* consumed, char* errmsg,
								 size_t errmsg_cap);

/* tools::Replayer::next loop (tools/replayer/src/replayer.cpp:56-98) over
 * in-memory samples: events older than the next image are delivered first
 * (FeatureDetector::updatePatches), then the image (FeatureDetector::newImage,
 * detector and LK on the device); ends when the events or the images run out.
 * frames: n_frames x height x width uint8.  Events must be in time order. */
int ebt_replay(ebt_ctx* ctx, const ebt_event* events, int64_t n_events, const uint8_t* frames,
			   const int64_t* frame_ts, int n_frames, int64_t* n_events_delivered, int64_t* n_optimizations);

/* ---- batched entry: F independent feature-patch problems ----------------
 * One Optimizer::optimize(Patch&) (optimizer.cpp:62-206) per problem.
 * ebt_frames_add registers a template frame (Optimizer::setGrad,
 * optimizer.cpp:15-30) and returns its slot (>= 0). */
int ebt_frames_add(ebt_ctx* ctx, const uint8_t* gray);
int ebt_frames_get(ebt_ctx* ctx, int slot, double* gx, double* gy);
int ebt_optimize_batch(ebt_ctx* ctx, const ebt_problem* problems, int64_t n_problems,
					   const ebt_event* events, int64_t n_events, ebt_solution* solutions,
					   int32_t* nabla /* n_problems x 625 or NULL */);
/* Device-resident variant: all pointers are device pointers on ctx's device;
 * asynchronous on the context stream (call ebt_sync to wait). */
int ebt_optimize_batch_device(ebt_ctx* ctx, const ebt_problem* d_problems, int64_t n_problems,
							  const ebt_event* d_events, ebt_solution* d_solutions,
							  int32_t* d_nabla);
int ebt_sync(ebt_ctx* ctx);

/* FeatureDetector::setParams -> reset() (feature_detector.cpp:734-742): drop all
 * patches, frames and records, keep the allocations. */
int ebt_reset(ebt_ctx* ctx);

/* Measured non-tensor FMA throughput of ctx's device in TFLOP/s (roofline
 * denominator of the LM kernels; not part of the reference's API). */
int ebt_measure_fma_peak(ebt_ctx* ctx, double* fp64_tflops, double* fp32_tflops);

/* Device-time of the kernels launched since ebt_timing_reset, measured with
 * CUDA events on the context stream: milliseconds and launch count per kernel
 * (0 stream kernel, 1 batched kernel, 2 gradient kernel, 3 window-size kernel,
 * 4 corner detector, 5 pyramid + Scharr, 6 LK tracker, 7 event-text parser). */
int ebt_timing_enable(ebt_ctx* ctx, int on);
int ebt_timing_get(ebt_ctx* ctx, int kernel, double* ms_total, int64_t* launches);
int ebt_timing_reset(ebt_ctx* ctx);

/* Profiling aid: cycle counters of the solver's leader warp, summed over all
 * solves since the last call (reset on read).  out8[0..4] = cycles spent in
 * request hand-off, the pixel pass, waiting for the other warps' partial sums,
 * totals + finalisation, and the rest of a trust-region iteration; out8[5] =
 * evaluations counted.  All zero unless the library was built with
 * -DEBT_PHASE_CLOCKS (the shipped build is not). */
int ebt_debug_phase_clocks(ebt_ctx* ctx, uint64_t* out8);

/* Diagnostics (not part of the reference's API): per live patch, what its CTA
 * did in the previous ebt_push_events: optimisations run, %globaltimer (ns) at
 * the CTA's entry and exit, SM id.  Used by tools/load_profile.py to show how
 * the stream kernel's duration follows the heaviest feature's serial chain. */
int ebt_debug_feature_load(ebt_ctx* ctx, int32_t* track_ids, int32_t* n_opt, uint64_t* t0_ns, uint64_t* t1_ns,
						   int32_t* sm, int cap, int* n_out);

/* Tuning knob of ebt_push_events (not part of the reference's API).  Before the
 * stream kernel runs, a prefilter marks, per feature, the events whose pixel
 * lies within `pixels` of the feature's patch rect as it stands at the push;
 * the feature then only visits marked events.  A rect that drifts further than
 * `pixels` during the push re-marks its remaining events, so results do not
 * depend on the value (0 forces a re-mark after nearly every optimisation).
 * Default 8. */
int ebt_set_prefilter_margin(ebt_ctx* ctx, double pixels);

/* Tuning knob of ebt_push_events (not part of the reference's API): which stream kernel runs.
 *   0            one CTA per feature (a leader warp + four pixel warps each)
 *   1 (default)  automatic: persistent CTAs of two leader warps sharing four pixel warps,
 *                features pulled from a queue, when more than two features per SM are live
 *                (or several streams are pushed at once) and the patch products and the cost
 *                map are off; one CTA per feature otherwise
 *   2            the persistent-CTA kernel whenever products and cost map are off
 * Same results either way (tests/test_gpu_stream.py runs both). */
int ebt_set_stream_kernel(ebt_ctx* ctx, int mode);

/* ABI self-description for binding generators/tests. */
size_t ebt_sizeof_event(void);
size_t ebt_sizeof_track(void);
size_t ebt_sizeof_opt_record(void);
size_t ebt_sizeof_problem(void);
size_t ebt_sizeof_solution(void);
size_t ebt_sizeof_params(void);
size_t ebt_sizeof_frontend_params(void);
size_t ebt_sizeof_camera(void);
int ebt_abi_version(void);

#ifdef __cplusplus
}
#endif
#endif /* EBT_H_ */
