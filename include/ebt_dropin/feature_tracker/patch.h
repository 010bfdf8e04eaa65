// Drop-in for the reference's <feature_tracker/patch.h>
// (/root/reference/implementation/feature_tracker/include/feature_tracker/patch.h:1-136).
//
// Put <repo>/include/ebt_dropin (and <repo>/include) in front of the reference's
// feature_tracker/include on the include path: the reference's callers —
// visual_odometry::Keyframe (vo/src/keyframe.cpp:5-14), the result writers and the
// viewer (evaluator.cpp:125-225, visualizer.cpp:158-217) — then compile unmodified
// against this tracker::Patch, which is a host-side SNAPSHOT of a patch whose life
// (event window, integration, optimisation) happens on the GPU.  Same names, same
// signatures, same types (cv::, Sophus, std::chrono through the reference's own
// <common/data_types.h>); the getters return what the reference's members would hold.
//
// Not carried over: the mutators that ARE the hot path (addEvent, integrateEvents,
// warpImage, updatePatchRect, ...): they run in the CUDA kernels and have no
// host-side meaning on a snapshot.
#pragma once

#include <common/data_types.h>

// the reference's patch.h brings these along and its callers (visual_odometry.h) rely on it
#include <spdlog/sinks/stdout_sinks.h>
#include <spdlog/spdlog.h>

#include <list>
#include <map>
#include <memory>
#include <optional>
#include <unordered_map>
#include <vector>

#include "ebt.h"

namespace tracker
{
using Corner = cv::Point2d;
using Corners = std::vector<Corner>;
using TrackId = int32_t;

class Patch
{
   public:
	// Patch::Patch + init (patch.cpp:8-35): a fresh patch around `corner`
	Patch(const Corner& corner, int extent, const common::timestamp_t& timestamp)
		: currentTimestamp_(timestamp), initTime_(timestamp)
	{
		patch_ = cv::Rect2d(corner.x - extent, corner.y - extent, 2 * extent + 1, 2 * extent + 1);
		initPoint_ = toCorner();
		integratedNabla_ = cv::Mat::zeros(static_cast<int>(patch_.height), static_cast<int>(patch_.width), CV_64F);
		motionCompensatedIntegratedNabla_ = integratedNabla_.clone();
		predictedNabla_ = integratedNabla_.clone();
		costMap_ = cv::Mat::zeros(1, 1, CV_64F);
		trajectory_.emplace_back(toCorner(), currentTimestamp_);
	}

	Corner toCorner() const { return Corner(patch_.x + (patch_.width - 1) / 2., patch_.y + (patch_.height - 1) / 2.); }
	bool isInPatch(const common::Point2i& point) const { return patch_.contains(point); }
	bool isLost() const { return lost_; }

	cv::Mat const& getIntegratedNabla() const { return integratedNabla_; }
	cv::Mat const& getPredictedNabla() const { return predictedNabla_; }
	cv::Mat const& getCompenatedIntegratedNabla() const { return motionCompensatedIntegratedNabla_; }
	cv::Rect2d const& getPatch() const { return patch_; }
	TrackId getTrackId() const { return trackId_; }
	const common::Pose2d& getWarp() const { return warp_; }
	float getFlow() const { return static_cast<float>(flowDir_); }
	const cv::Mat& getCostMap() const { return costMap_; }
	bool isInit() { return init_; }
	std::vector<common::Sample<common::Point2d>> const& getTrajectory() const { return trajectory_; }
	size_t getNumOfEvents() const { return numOfEvents_; }
	cv::Rect2d getInitPatch() const
	{
		return cv::Rect2d(initPoint_.x - (patch_.width - 1) / 2., initPoint_.y - (patch_.height - 1) / 2., patch_.width,
						  patch_.height);
	}
	common::timestamp_t getCurrentTimestamp() const { return currentTimestamp_; }
	common::timestamp_t getTimeWithoutUpdate() const { return timeWithoutUpdate_; }
	const std::vector<double>& getFinalCosts() const { return finalCosts_; }
	common::timestamp_t getTimeLastUpdate() const { return timeLastUpdate_; }
	common::timestamp_t getInitTime() const { return initTime_; }

	void setLost() { lost_ = true; }
	void setTrackId(TrackId trackId) { trackId_ = trackId; }
	void setTs(const common::timestamp_t& ts) { currentTimestamp_ = ts; }

   private:
	friend class FeatureDetector;
	bool init_ = false;
	bool lost_ = false;
	cv::Rect2d patch_;
	common::Point2d initPoint_;
	TrackId trackId_ = -1;
	common::timestamp_t currentTimestamp_{0};
	common::timestamp_t timeLastUpdate_{0};
	common::timestamp_t timeWithoutUpdate_{static_cast<int64_t>(1e7)};
	common::timestamp_t initTime_{0};
	size_t numOfEvents_ = 75;
	cv::Mat integratedNabla_;
	cv::Mat motionCompensatedIntegratedNabla_;
	cv::Mat predictedNabla_;
	cv::Mat costMap_;
	double flowDir_ = 0;
	common::Pose2d warp_;
	std::vector<common::Sample<common::Point2d>> trajectory_;
	std::vector<double> finalCosts_;
};

using Patches = std::list<Patch>;

}  // namespace tracker
