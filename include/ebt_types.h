/* ebt_types.h — plain-C records shared by the C-ABI (include/ebt.h), the CPU
 * oracle (oracle/) and the CUDA implementation (event_based_odomety_b200/csrc).
 *
 * Vocabulary follows the reference:
 *   common::EventSample   /root/reference/common/include/common/data_types.h:12-38
 *   tracker::Patch state  /root/reference/implementation/feature_tracker/include/feature_tracker/patch.h:98-130
 */
#ifndef EBT_TYPES_H_
#define EBT_TYPES_H_

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define EBT_PATCH_EXTENT 12
#define EBT_PATCH_SIDE 25
#define EBT_PATCH_PIXELS 625
#define EBT_MAX_WINDOW 300 /* Patch::maxNumOfEvents_, patch.h:114 */
#define EBT_COST_HISTORY 8 /* >= 6 newest final costs (optimizer.cpp:138-157) */

/* One event, 16 bytes = one 128-bit load.  Replaces common::EventSample
 * (24 B: cv::Point2i + EventPolarity + std::chrono::microseconds). */
typedef struct ebt_event
{
	int64_t t_us;    /* EventSample::timestamp.count() */
	uint16_t x;      /* EventSample::value.point.x */
	uint16_t y;      /* EventSample::value.point.y */
	int8_t polarity; /* EventSample::value.sign: -1 or +1 */
	int8_t pad_[3];
} ebt_event;

/* Snapshot of one tracker::Patch (the scalars of patch.h:98-130). */
typedef struct ebt_track
{
	int32_t track_id;       /* Patch::trackId_ */
	int32_t lost;           /* Patch::lost_ */
	int32_t init;           /* Patch::init_ (flow/warp initialised) */
	int32_t num_of_events;  /* Patch::numOfEvents_ (window size) */
	int32_t counter;        /* Patch::counter_ (events since last optimise) */
	int32_t window_size;    /* Patch::events_.size() */
	int32_t n_optimizations;/* Patch::finalCosts_.size() */
	int32_t n_trajectory;   /* Patch::trajectory_.size() */
	double rect_x, rect_y;  /* Patch::patch_.tl() (25x25) */
	double init_x, init_y;  /* Patch::initPoint_ */
	double warp[4];         /* Sophus::SE2d::data(): cos, sin, tx, ty */
	double flow_dir;        /* Patch::flowDir_ */
	double last_cost;       /* Patch::finalCosts_.back() or 0 */
	int64_t current_ts;     /* Patch::currentTimestamp_ */
	int64_t time_last_update;    /* Patch::timeLastUpdate_ */
	int64_t time_without_update; /* Patch::timeWithoutUpdate_ */
	int64_t init_time;      /* Patch::initTime_ (template frame) */
} ebt_track;

/* One completed Optimizer::optimize (optimizer.cpp:62-206): what the call
 * appended to Patch::trajectory_ / Patch::finalCosts_ /
 * Optimizer::vectorFinalCost_, plus solver diagnostics. */
typedef struct ebt_opt_record
{
	int32_t track_id;
	int32_t n_successful;   /* accepted LM steps (Summary::num_successful_steps-1) */
	int32_t n_unsuccessful; /* rejected + invalid LM steps */
	int32_t termination;    /* 0 NO_CONVERGENCE, 1 CONVERGENCE, 2 FAILURE */
	int32_t lost;           /* patch lost by this call (cost filter / border) */
	int32_t num_of_events;  /* window size chosen after the call */
	int64_t ts;             /* Patch::currentTimestamp_ after integrateEvents */
	int64_t event_index;    /* index in the pushed stream of the triggering event */
	double x, y;            /* Patch::toCorner() after the call */
	double final_cost;      /* Summary::final_cost */
	double warp[4];
	double flow_dir;
} ebt_opt_record;

/* One independent feature-patch problem for the batched solve entry point:
 * the state a Patch has when Optimizer::optimize(patch) is entered. */
typedef struct ebt_problem
{
	double init_x, init_y; /* Patch::initPoint_ */
	double warp[4];        /* cos, sin, tx, ty */
	double flow_dir;       /* Patch::flowDir_ (read back as float, patch.h:54) */
	int32_t n_events;      /* events in the window, newest first */
	int32_t event_offset;  /* first event of this window in the events array */
	int32_t frame_slot;    /* gradient grid (template frame) to use */
	int32_t pad_;
} ebt_problem;

typedef struct ebt_solution
{
	double warp[4];
	double flow_dir;        /* after fmod(flow, 2*pi) */
	double final_cost;
	double x, y;            /* new patch centre */
	int32_t n_successful;
	int32_t n_unsuccessful;
	int32_t termination;
	int32_t n_passes;       /* pixel passes of the solve: 100 * (cost+Jacobian passes) + cost-only passes; -1 = not reported (CPU checkers) */
	int64_t ts;             /* int32((t_front+t_back)*0.5) */
} ebt_solution;

/* common::CameraModelParams<double> (common/include/common/camera_model.h:12-24),
 * same field order (CameraModel::fromData reinterprets a double[9]). */
typedef struct ebt_camera
{
	double fx, fy, cx, cy, k1, k2, k3, p1, p2;
} ebt_camera;

#ifdef __cplusplus
}
#endif
#endif /* EBT_TYPES_H_ */
