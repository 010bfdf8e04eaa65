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
	bool drawCostMap = false;
	int maxNumIterations = 10;
	int numThreads = 1;  // unused: one warp or CTA per feature on the device
	double optimizerThreshold = 0.6;
	double huberLoss = 0.3;
	int costMapWidth = 11;   // fixed at 11 x 11 on the device
	int costMapHeight = 11;
	// seconds in double to microseconds
	double patchTimeWithoutUpdateScale = 1e6;
};

struct OptimizerFinalLoss
{
	tracker::TrackId trackId;
	double lossValue;
	int64_t timeStampMicrosecond;
};
}  // namespace tracker
