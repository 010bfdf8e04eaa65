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
	// What a snapshot holds: the tracker's state of one patch as the C-ABI reports it (ebt_track,
	// ebt_get_patch, ebt_get_trajectory, ebt_get_final_costs), in the reference's types.  The
	// drop-in FeatureDetector fills it; the reference's callers only see the getters below.
	struct View
	{
		cv::Rect2d rect;                  // 25x25 window in image coordinates (top-left, size)
		common::Point2d origin;           // where the patch was created
		TrackId id = -1;
		bool has_flow = false;            // LK flow / warp initialised (Patch::init_)
		bool lost = false;
		size_t window = 75;               // events per optimisation window (patch.cpp:22)
		double flow_angle = 0;            // rad
		common::Pose2d warp;
		common::timestamp_t now{0}, last_update{0}, created{0};
		common::timestamp_t stale_after{static_cast<int64_t>(1e7)};
		cv::Mat events_patch;             // integrated event polarities
		cv::Mat compensated_patch;        // the same, motion compensated (getter-only product)
		cv::Mat predicted_patch;          // template gradient under the current warp and flow (getter-only product)
		cv::Mat cost_map;                 // 11x11 sweep, when OptimizerParams::drawCostMap is on
		std::vector<common::Sample<common::Point2d>> path;  // image-time samples and kept optimisations
		std::vector<double> costs;        // final cost of every optimisation
	};

	// Patch::Patch + init (patch.cpp:8-35): a fresh patch of half-width `extent` around `corner`
	Patch(const Corner& corner, int extent, const common::timestamp_t& timestamp)
	{
		const double side = 2 * extent + 1;
		v_.rect = cv::Rect2d(corner.x - extent, corner.y - extent, side, side);
		v_.now = v_.created = timestamp;
		v_.origin = toCorner();
		v_.events_patch = cv::Mat::zeros(static_cast<int>(side), static_cast<int>(side), CV_64F);
		v_.compensated_patch = v_.events_patch.clone();
		v_.predicted_patch = v_.events_patch.clone();
		v_.cost_map = cv::Mat::zeros(1, 1, CV_64F);
		v_.path.emplace_back(v_.origin, timestamp);
	}

	// ---- geometry (patch.cpp:161-176)
	Corner toCorner() const { return Corner(v_.rect.x + (v_.rect.width - 1) / 2., v_.rect.y + (v_.rect.height - 1) / 2.); }
	cv::Rect2d const& getPatch() const { return v_.rect; }
	cv::Rect2d getInitPatch() const
	{
		const double hw = (v_.rect.width - 1) / 2., hh = (v_.rect.height - 1) / 2.;
		return cv::Rect2d(v_.origin.x - hw, v_.origin.y - hh, v_.rect.width, v_.rect.height);
	}
	bool isInPatch(const common::Point2i& point) const { return v_.rect.contains(point); }

	// ---- identity and life cycle
	TrackId getTrackId() const { return v_.id; }
	void setTrackId(TrackId trackId) { v_.id = trackId; }
	bool isInit() { return v_.has_flow; }
	bool isLost() const { return v_.lost; }
	void setLost() { v_.lost = true; }
	common::timestamp_t getInitTime() const { return v_.created; }
	common::timestamp_t getCurrentTimestamp() const { return v_.now; }
	void setTs(const common::timestamp_t& ts) { v_.now = ts; }
	common::timestamp_t getTimeLastUpdate() const { return v_.last_update; }
	common::timestamp_t getTimeWithoutUpdate() const { return v_.stale_after; }

	// ---- what the optimiser estimated (the reference reads the flow back as a float, patch.h:54)
	const common::Pose2d& getWarp() const { return v_.warp; }
	float getFlow() const { return static_cast<float>(v_.flow_angle); }
	size_t getNumOfEvents() const { return v_.window; }
	const std::vector<double>& getFinalCosts() const { return v_.costs; }
	std::vector<common::Sample<common::Point2d>> const& getTrajectory() const { return v_.path; }

	// ---- 25x25 images (CV_64F) and the cost map
	cv::Mat const& getIntegratedNabla() const { return v_.events_patch; }
	cv::Mat const& getCompenatedIntegratedNabla() const { return v_.compensated_patch; }
	cv::Mat const& getPredictedNabla() const { return v_.predicted_patch; }
	const cv::Mat& getCostMap() const { return v_.cost_map; }

   private:
	friend class FeatureDetector;
	View v_;
};

using Patches = std::list<Patch>;

}  // namespace tracker
