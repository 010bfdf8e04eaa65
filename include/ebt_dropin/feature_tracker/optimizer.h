// Drop-in for the reference's <feature_tracker/optimizer.h>
// (/root/reference/implementation/feature_tracker/include/feature_tracker/optimizer.h:14-34):
// the two structs its callers use.  tracker::Optimizer itself (Ceres problem per
// patch, optimizer.cpp:62-206) is the LM kernel on the device and has no host class.
#pragma once

#include <cstdint>

#include "feature_tracker/patch.h"

namespace tracker
{
struct OptimizerParams
{
	bool drawCostMap = false;          // ebt_set_cost_map: 121 extra pixel passes per optimisation
	int maxNumIterations = 10;         // ebt_params::max_iterations (Ceres max_num_iterations)
	int numThreads = 1;                // unused: one warp or CTA per feature on the device
	double optimizerThreshold = 0.6;   // ebt_params::optimizer_threshold (median-of-5 cost filter)
	double huberLoss = 0.3;            // ebt_params::huber_loss
	int costMapWidth = 11;             // fixed at 11 x 11 on the device
	int costMapHeight = 11;
	double patchTimeWithoutUpdateScale = 1e6;  // seconds in double to microseconds (fixed on the device)
};

struct OptimizerFinalLoss  // one per optimisation: the (track, cost, time) projection of ebt_opt_record
{
	tracker::TrackId trackId;      // ebt_opt_record::track_id
	double lossValue;              // ebt_opt_record::final_cost
	int64_t timeStampMicrosecond;  // ebt_opt_record::ts
};
}  // namespace tracker
