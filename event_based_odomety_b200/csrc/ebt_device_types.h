// ebt_device_types.h — records shared by the host side of the C-ABI
// (ebt_api.cu) and the sm_100a kernels (ebt_kernels.cu).
#pragma once

#include <stdint.h>

#include "../../include/ebt_types.h"

// live template frames (FeatureDetector::optimizers_) a context starts with room for; the
// table grows on demand (the reference's map is unbounded)
#define EBT_INIT_FRAMES 64
#ifndef EBT_THREADS
#define EBT_THREADS 128     // threads per feature CTA
#endif
// per-feature window ring capacity: a power of two >= 300 + one chunk of EBT_THREADS
// resident feature CTAs per SM the stream/batched kernels are built for (Makefile MIN_CTAS)
#ifndef EBT_MIN_CTAS
#define EBT_MIN_CTAS 3
#endif
#define EBT_RESIDENT_CTAS EBT_MIN_CTAS
#define EBT_RING (EBT_THREADS <= 212 ? 512 : 1024)

// Device mirror of the scalars of tracker::Patch
// (/root/reference/implementation/feature_tracker/include/feature_tracker/patch.h:98-130).
// The event window (Patch::events_) lives in rings[store_slot][EBT_RING]
// (element k-th oldest = ring[(win_tail + k) & (EBT_RING-1)]), the integrated
// patch (Patch::integratedNabla_) in nabla[store_slot][625] as int32.
struct DevPatch
{
	double rx, ry;   // patch_.tl()
	double ix, iy;   // initPoint_
	double warp[4];  // Sophus::SE2d data: cos sin tx ty
	double flow;     // flowDir_
	double costs[EBT_COST_HISTORY];  // newest finalCosts_, ring indexed by n_costs
	long long cur_ts, time_last_update, time_without_update, init_time;
	int init, lost, track_id, num_of_events;
	int counter, win_tail, win_size, n_costs;
	int frame_slot, store_slot;
	int last_n_opt;  // optimisations in the previous push (CTA ordering estimate)
	int n_traj;      // Patch::trajectory_.size(): image-time samples (host) + kept optimisations (device)
	// the two newest entries of Patch::trajectory_ (what integrateMotionCompensatedEvents reads,
	// patch.cpp:91-92): [0] = second newest, [1] = newest
	double traj_x[2], traj_y[2];
	long long traj_ts[2];
	// load diagnostics of the previous push (ebt_debug_feature_load): %globaltimer at the
	// CTA's entry and exit, and the SM it ran on
	unsigned long long last_t0_ns, last_t1_ns;
	int last_sm;
	// feature-batch sharding of one stream over several GPUs (ebt_set_shard): this patch is
	// tracked by another shard; it stays in the list (association and list order are global)
	// but no kernel of this context touches it
	int foreign;
};

// Solver constants: the reference's OptimizerParams
// (feature_tracker/optimizer.h:16-27) + the Ceres 2.0 defaults its Solve() call
// inherits (thirdparty/ceres-solver/include/ceres/solver.h:264-322).
struct DevParams
{
	int width, height;
	int max_iterations;           // 10
	int init_num_events;          // 75  (DetectorParams::initNumEvents)
	int min_num_events;           // 100 (Patch::minNumOfEvents_)
	int max_num_events;           // 300 (Patch::maxNumOfEvents_)
	int batch_stride;             // 30  (Patch::isReady)
	int border_margin;            // 5   (updateNumOfEvents)
	double huber_a;               // 0.3
	double optimizer_threshold;   // 0.6
	double initial_radius;        // 1e4
	double max_radius;            // 1e16
	double min_radius;            // 1e-32
	double min_relative_decrease; // 1e-3
	double min_lm_diagonal;       // 1e-6
	double max_lm_diagonal;       // 1e32
	double function_tolerance;    // 1e-6
	double gradient_tolerance;    // 1e-10
	double parameter_tolerance;   // 1e-8
	int max_consecutive_nonmonotonic_steps;  // 5
	int max_consecutive_invalid_steps;       // 5
	// Getter-only products of tracker::Patch (ebt_set_products): 0 none, 1 after every
	// optimisation (what the reference always does), 2 also at image time (drawImages).
	int products;
	int pad_;
	int* mc_nablas;       // [store_slot][625] Patch::motionCompensatedIntegratedNabla_
	double* pred_nablas;  // [store_slot][625] Patch::predictedNabla_
	// OptimizerParams::drawCostMap (ebt_set_cost_map): null = off; stream kernel
	// [store_slot][121], batched kernel [problem][121] = Patch::costMap_ (11x11)
	double* cost_maps;
};

// Gradient grids of the live template frames: interleaved (gx, gy) doubles,
// row stride = width (Optimizer::setGrad, optimizer.cpp:15-30).  `grid` is a
// device array of grid pointers indexed by DevPatch::frame_slot / ebt_problem::
// frame_slot; the host grows it as frames are added.
struct FrameTable
{
	const double2* const* grid;
	int latest;  // slot FeatureDetector::gradX_/gradY_ refer to
	int pad;
};

// One stream of a multi-stream push (ebt_launch_track_multi): where its features,
// windows, event batch, prefilter rows, template grids and records live.
struct StreamDesc
{
	DevPatch* patches;
	ebt_event* rings;
	int* nablas;
	const ebt_event* events;   // this push's batch, stream order
	long long n_events;
	long long event_base;      // stream index of events[0] (ebt_opt_record::event_index)
	unsigned* mask;            // n_patches x words_per_row prefilter words
	const double2* const* grids;  // FrameTable::grid
	ebt_opt_record* records;
	unsigned long long* n_records;
	unsigned long long record_cap;
	double margin;             // prefilter margin, pixels
	int n_patches;
	int words_per_row;
	int latest;                // FrameTable::latest
	int pad_;
};

// One unit of work of the multi-stream kernel: feature `feature` of stream `stream`.
struct WorkItem
{
	int stream, feature;
};
