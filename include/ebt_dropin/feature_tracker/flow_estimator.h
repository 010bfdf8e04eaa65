// Drop-in for the reference's <feature_tracker/flow_estimator.h>
// (/root/reference/implementation/feature_tracker/include/feature_tracker/flow_estimator.h:11-17):
// the parameter struct.  tracker::FlowEstimator (one cv::calcOpticalFlowPyrLK call per
// uninitialised patch, flow_estimator.cpp:27-103) runs inside FeatureDetector::newImage
// on the device (csrc/ebt_frontend.cu: pyramid, Scharr, LK in one launch for all points).
#pragma once

#include "feature_tracker/patch.h"

namespace tracker
{
struct FlowEstimatorParams
{
	int32_t numLevels = 3;                     // ebt_frontend_params::lk_levels (maxLevel of calcOpticalFlowPyrLK)
	cv::Size window = {21, 21};                // fixed on the device (lk_track_kernel's 21 x 21 template)
	double patchTimeWithoutUpdateScale = 1e6;  // seconds in double to microseconds (fixed on the device)
};
}  // namespace tracker
