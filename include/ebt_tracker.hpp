// ebt_tracker.hpp — C++ facade over the C-ABI (include/ebt.h) with the reference's
// class and method names, so that the callers of tracker::FeatureDetector /
// tracker::Patch (tools/evaluator/src/evaluator.cpp:32-101, visual_odometry/src/
// keyframe.cpp:9-13, result writers evaluator.cpp:125-225) keep compiling with
// one include swapped.  Header-only; link with libebt_cuda.so.
//
// Differences a caller sees (all forced by the missing OpenCV / batching):
//  * images are `const uint8_t*` (CV_8U, row-major) instead of cv::Mat;
//  * the Harris detector and the LK tracker the reference calls inside newImage
//    (feature_detector.cpp:568-583, flow_estimator.cpp:84-103) run on the device
//    (ebt_new_image); a host that wants to keep cv::goodFeaturesToTrack /
//    calcOpticalFlowPyrLK installs them as std::function hooks (both or neither);
//  * updatePatches(event) buffers; the batch is flushed on newImage, on any
//    getter, or when `flushThreshold` events are pending.  Results equal the
//    per-event loop.
//  * Patch objects are snapshots (value semantics, like the reference's copies).
#pragma once

#include <cstdint>
#include <functional>
#include <list>
#include <stdexcept>
#include <string>
#include <vector>

#include "ebt.h"

namespace tracker
{
struct Corner
{
	double x = 0, y = 0;
};
using Corners = std::vector<Corner>;
using TrackId = int32_t;

struct Rect2d
{
	double x = 0, y = 0, width = 0, height = 0;
};

// common::Sample<common::Point2d> (common/include/common/data_types.h:12-24)
struct TrajectorySample
{
	Corner value;
	int64_t timestamp = 0;  // microseconds
};

// common::EventSample (data_types.h:26-38)
struct EventSample
{
	int x = 0, y = 0;
	int sign = 1;           // EventPolarity: -1 / +1
	int64_t timestamp = 0;  // microseconds
};

struct OptimizerParams  // feature_tracker/optimizer.h:16-27
{
	bool drawCostMap = false;  // 121 extra pixel passes per optimisation (optimizer.cpp:191-197)
	int maxNumIterations = 10;
	double optimizerThreshold = 0.6;
	double huberLoss = 0.3;
};

struct DetectorParams  // feature_tracker/feature_detector.h:10-31 (hot-path fields)
{
	double associationDistance = 5;
	int32_t patchExtent = 12;  // fixed: 25x25 patches
	int imageWidth = 240, imageHeight = 180;
	OptimizerParams optimizerParams = {};
	int initNumEvents = 75;
	size_t maxPatches = 100;
	size_t flushThreshold = 1 << 20;  // events buffered before an automatic flush
};

struct OptimizerFinalLoss  // feature_tracker/optimizer.h:29-34
{
	TrackId trackId;
	double lossValue;
	int64_t timeStampMicrosecond;
};

// Snapshot of tracker::Patch (feature_tracker/patch.h:14-130): the getters the
// reference's callers use.
class Patch
{
   public:
	Corner toCorner() const { return {t_.rect_x + 12.0, t_.rect_y + 12.0}; }
	TrackId getTrackId() const { return t_.track_id; }
	bool isLost() const { return t_.lost != 0; }
	bool isInit() const { return t_.init != 0; }
	Rect2d getPatch() const { return {t_.rect_x, t_.rect_y, 25, 25}; }
	Rect2d getInitPatch() const { return {t_.init_x - 12.0, t_.init_y - 12.0, 25, 25}; }
	float getFlow() const { return static_cast<float>(t_.flow_dir); }
	const double* getWarp() const { return t_.warp; }  // Sophus::SE2d::data(): cos sin tx ty
	size_t getNumOfEvents() const { return static_cast<size_t>(t_.num_of_events); }
	int64_t getCurrentTimestamp() const { return t_.current_ts; }
	int64_t getTimeLastUpdate() const { return t_.time_last_update; }
	int64_t getTimeWithoutUpdate() const { return t_.time_without_update; }
	int64_t getInitTime() const { return t_.init_time; }
	const std::vector<TrajectorySample>& getTrajectory() const { return trajectory_; }
	const std::vector<double>& getFinalCosts() const { return finalCosts_; }
	const std::vector<double>& getIntegratedNabla() const { return nabla_; }  // 25x25 row-major
	// 11x11 row-major once optimised with OptimizerParams::drawCostMap, else a single 0 (patch.cpp:29)
	const std::vector<double>& getCostMap() const { return costMap_; }

   private:
	friend class FeatureDetector;
	ebt_track t_{};
	std::vector<TrajectorySample> trajectory_;
	std::vector<double> finalCosts_;
	std::vector<double> nabla_;
	std::vector<double> costMap_ = std::vector<double>(1, 0.0);
};
using Patches = std::list<Patch>;

// tracker::FeatureDetector (feature_tracker/feature_detector.h:33-120), tracking part.
class FeatureDetector
{
   public:
	using CornerDetector = std::function<Corners(const uint8_t* gray, int width, int height)>;
	// LK from the previous to the current image at `point`; false = status 0
	using FlowTracker = std::function<bool(const Corner& point, float& nextX, float& nextY)>;

	explicit FeatureDetector(const DetectorParams& params, int device = 0) : params_(params)
	{
		ebt_params p;
		check(ebt_default_params(&p, params.imageWidth, params.imageHeight), "ebt_default_params");
		p.max_patches = static_cast<int32_t>(params.maxPatches);
		p.init_num_events = params.initNumEvents;
		p.association_distance = params.associationDistance;
		p.max_iterations = params.optimizerParams.maxNumIterations;
		p.huber_loss = params.optimizerParams.huberLoss;
		p.optimizer_threshold = params.optimizerParams.optimizerThreshold;
		if (ebt_create(&p, device, &ctx_) != EBT_OK)
			throw std::runtime_error(std::string("ebt_create: ") + ebt_last_error(nullptr));
		if (params.optimizerParams.drawCostMap) check(ebt_set_cost_map(ctx_, 1), "ebt_set_cost_map");
	}
	~FeatureDetector()
	{
		if (ctx_) ebt_destroy(ctx_);
	}
	FeatureDetector(const FeatureDetector&) = delete;
	FeatureDetector& operator=(const FeatureDetector&) = delete;

	void setCornerDetector(CornerDetector f) { detect_ = std::move(f); }
	void setFlowTracker(FlowTracker f) { flow_ = std::move(f); }

	// FeatureDetector::newImage (feature_detector.cpp:493-541)
	void newImage(const uint8_t* gray, int64_t timestampUs)
	{
		flush();
		if (!detect_ && !flow_)
		{
			check(ebt_new_image(ctx_, gray, timestampUs), "ebt_new_image");
			int nc = 0;
			check(ebt_get_features(ctx_, nullptr, 0, &nc), "ebt_get_features");
			std::vector<double> cxy(static_cast<size_t>(nc) * 2);
			if (nc) check(ebt_get_features(ctx_, cxy.data(), nc, &nc), "ebt_get_features");
			corners_.clear();
			for (int i = 0; i < nc; ++i) corners_.push_back(Corner{cxy[2 * i], cxy[2 * i + 1]});
			dirty_ = true;
			return;
		}
		corners_ = detect_ ? detect_(gray, params_.imageWidth, params_.imageHeight) : Corners();
		std::vector<double> xy;
		for (const Corner& c : corners_)
		{
			xy.push_back(c.x);
			xy.push_back(c.y);
		}
		check(ebt_new_frame_begin(ctx_, gray, timestampUs, xy.data(), static_cast<int>(corners_.size())),
			  "ebt_new_frame_begin");
		int n = 0;
		check(ebt_get_uninit(ctx_, nullptr, nullptr, 0, &n), "ebt_get_uninit");
		std::vector<int32_t> ids(static_cast<size_t>(n));
		std::vector<double> pts(static_cast<size_t>(n) * 2);
		if (n) check(ebt_get_uninit(ctx_, ids.data(), pts.data(), n, &n), "ebt_get_uninit");
		std::vector<int32_t> found(static_cast<size_t>(n));
		std::vector<float> next(static_cast<size_t>(n) * 2);
		for (int i = 0; i < n; ++i)
		{
			float nx = 0, ny = 0;
			found[i] = (flow_ && flow_(Corner{pts[2 * i], pts[2 * i + 1]}, nx, ny)) ? 1 : 0;
			next[2 * i] = nx;
			next[2 * i + 1] = ny;
		}
		if (n) check(ebt_set_flows(ctx_, ids.data(), found.data(), next.data(), n), "ebt_set_flows");
		check(ebt_new_frame_end(ctx_), "ebt_new_frame_end");
		dirty_ = true;
	}

	// FeatureDetector::updatePatches (feature_detector.cpp:585-619): buffered
	void updatePatches(const EventSample& e)
	{
		ebt_event ev{};
		ev.t_us = e.timestamp;
		ev.x = static_cast<uint16_t>(e.x);
		ev.y = static_cast<uint16_t>(e.y);
		ev.polarity = static_cast<int8_t>(e.sign > 0 ? 1 : -1);
		pending_.push_back(ev);
		if (pending_.size() >= params_.flushThreshold) flush();
	}
	void updatePatches(const ebt_event* events, size_t n)
	{
		pending_.insert(pending_.end(), events, events + n);
		if (pending_.size() >= params_.flushThreshold) flush();
	}
	// addEvent only feeds the (out-of-scope) motion compensation in the reference
	void addEvent(const EventSample&) {}

	void flush()
	{
		if (pending_.empty()) return;
		check(ebt_push_events(ctx_, pending_.data(), static_cast<int64_t>(pending_.size()), nullptr),
			  "ebt_push_events");
		pending_.clear();
		dirty_ = true;
	}

	Patches const& getPatches()
	{
		refresh();
		return patches_;
	}
	Patches const& getArchivedPatches()
	{
		refresh();
		return archived_;
	}
	Corners const& getFeatures() const { return corners_; }

	// FeatureDetector::preExit (feature_detector.cpp:485-491): archive what is left
	void preExit()
	{
		refresh();
		for (const Patch& p : patches_) archived_.push_back(p);
	}

	// Optimizer::getFinalCosts via FeatureDetector::getOptimizedFinalCosts
	// (feature_detector.h:80-83): all optimisations, in execution order.
	std::vector<OptimizerFinalLoss> getOptimizedFinalCosts()
	{
		flush();
		int64_t n = 0;
		check(ebt_get_records(ctx_, nullptr, 0, &n), "ebt_get_records");
		std::vector<ebt_opt_record> r(static_cast<size_t>(n));
		if (n) check(ebt_get_records(ctx_, r.data(), n, &n), "ebt_get_records");
		std::vector<OptimizerFinalLoss> out;
		for (const auto& k : r) out.push_back({k.track_id, k.final_cost, k.ts});
		return out;
	}

	ebt_ctx* handle() { return ctx_; }

   private:
	void check(int rc, const char* what) const
	{
		if (rc < 0) throw std::runtime_error(std::string(what) + ": " + ebt_last_error(ctx_));
	}
	void fill(Patches& dst, int archived)
	{
		int n = 0;
		check(ebt_get_tracks(ctx_, nullptr, 0, archived, &n), "ebt_get_tracks");
		std::vector<ebt_track> t(static_cast<size_t>(n));
		if (n) check(ebt_get_tracks(ctx_, t.data(), n, archived, &n), "ebt_get_tracks");
		dst.clear();
		for (const ebt_track& k : t)
		{
			Patch p;
			p.t_ = k;
			std::vector<double> xyt(static_cast<size_t>(k.n_trajectory) * 3);
			int m = 0;
			check(ebt_get_trajectory(ctx_, k.track_id, xyt.data(), k.n_trajectory, &m), "ebt_get_trajectory");
			for (int i = 0; i < k.n_trajectory; ++i)
				p.trajectory_.push_back({{xyt[3 * i], xyt[3 * i + 1]}, static_cast<int64_t>(xyt[3 * i + 2])});
			p.finalCosts_.resize(static_cast<size_t>(k.n_optimizations));
			check(ebt_get_final_costs(ctx_, k.track_id, p.finalCosts_.data(), k.n_optimizations, &m),
				  "ebt_get_final_costs");
			p.nabla_.resize(EBT_PATCH_PIXELS);
			check(ebt_get_patch(ctx_, k.track_id, 0, p.nabla_.data()), "ebt_get_patch");
			if (params_.optimizerParams.drawCostMap && k.n_optimizations > 0)
			{
				p.costMap_.resize(121);
				check(ebt_get_cost_map(ctx_, k.track_id, p.costMap_.data()), "ebt_get_cost_map");
			}
			dst.push_back(std::move(p));
		}
	}
	void refresh()
	{
		flush();
		if (!dirty_) return;
		fill(patches_, 0);
		fill(archived_, 1);
		dirty_ = false;
	}

	DetectorParams params_;
	ebt_ctx* ctx_ = nullptr;
	CornerDetector detect_;
	FlowTracker flow_;
	std::vector<ebt_event> pending_;
	Patches patches_, archived_;
	Corners corners_;
	bool dirty_ = true;
};
}  // namespace tracker
