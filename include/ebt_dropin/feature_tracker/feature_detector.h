// Drop-in for the reference's <feature_tracker/feature_detector.h>
// (/root/reference/implementation/feature_tracker/include/feature_tracker/feature_detector.h:1-145):
// tracker::DetectorParams and tracker::FeatureDetector with the reference's own
// signatures, on top of the C-ABI (include/ebt.h, libebt_cuda.so).  With
// <repo>/include/ebt_dropin and <repo>/include in front of the reference's
// feature_tracker/include, tools::Evaluator (tools/evaluator/src/evaluator.cpp:
// 17-20, 32-44, 69-70, 85, 109-122) and visual_odometry (vo/src/keyframe.cpp:5-14)
// compile and run unmodified; link libebt_cuda.so instead of libfeature_tracker.
//
// What differs behind the same calls:
//  * updatePatches(event) buffers; the batch goes to the GPU at the next newImage,
//    at any getter, or when DetectorParams::flushThreshold events are pending.
//    Between two images a patch's evolution depends only on its own state and the
//    ordered event stream (feature_detector.cpp:587-617), so the results are those of
//    the per-event loop (tests/test_gpu_stream.py).
//  * detectFeatures + FlowEstimator (cv::goodFeaturesToTrack, cv::calcOpticalFlowPyrLK)
//    run on the device inside newImage.
//  * getPatches() returns snapshots (value semantics, as the reference's copies into
//    Keyframe / archivedPatches_).
//  * Event motion compensation (addEvent / compensateEventsContrast / integrateEvents /
//    get*EventImage: a separate algorithm that only feeds the GUI) is not on the hot
//    path: those calls are accepted and do nothing, the two images stay zero.
//  * Patch::getPredictedNabla / getCompenatedIntegratedNabla are filled when
//    DetectorParams::drawImages or ::patchProducts is set (they cost ~8 % and only
//    feed the viewer); getCostMap when optimizerParams.drawCostMap is set.
#pragma once

#include <common/data_types.h>

#include <cstring>
#include <list>
#include <stdexcept>
#include <string>
#include <vector>

#include "ebt.h"
#include "feature_tracker/flow_estimator.h"
#include "feature_tracker/optimizer.h"
#include "feature_tracker/patch.h"

namespace tracker
{
struct DetectorParams
{
	double qualityLevel = 0.01;
	double minDistance = 10;
	double associationDistance = 5;
	int32_t patchExtent = 12;  // fixed: 25 x 25 patches on the device
	int32_t blockSize = 3;     // fixed
	cv::Size imageSize = {240, 180};
	bool drawImages = false;
	OptimizerParams optimizerParams = {};
	int initNumEvents = 75;
	u_long maxNumEventsToStore = 15000;  // motion compensation: unused
	bool useAverageFlow = true;          // "
	bool optimizeFlowTV = true;          // "
	bool useL1 = false;                  // "
	cv::Size patchCompensateSize = {20, 20};
	double compensateTVweight = 1e3;
	double compensateTVHuberLoss = 10;
	double compensateScale = 1e-3;
	uint compensateMinNumEvents = 100;
	size_t maxPatches = 100;
	// ---- additions (defaults keep the reference's behaviour) ----
	int device = 0;                    // CUDA device of this detector
	size_t flushThreshold = 1 << 20;   // events buffered before an automatic flush
	bool patchProducts = false;        // fill getPredictedNabla / getCompenatedIntegratedNabla after every optimisation
};

class FeatureDetector
{
   public:
	FeatureDetector(const DetectorParams& params) : params_(params) { create(); }
	~FeatureDetector()
	{
		if (ctx_) ebt_destroy(ctx_);
	}
	FeatureDetector(const FeatureDetector&) = delete;
	FeatureDetector& operator=(const FeatureDetector&) = delete;

	// feature_detector.cpp:485-491
	void preExit()
	{
		refresh();
		for (const Patch& p : patches_) archivedPatches_.push_back(p);
	}

	// feature_detector.cpp:493-541; image.value: CV_8U, imageSize (3 channels are converted
	// when the real OpenCV is present)
	void newImage(const common::ImageSample& image)
	{
		flush();
		const cv::Mat gray = asGray(image.value);
		check(ebt_new_image(ctx_, gray.data, image.timestamp.count()), "ebt_new_image");
		readFeatures();
		dirty_ = true;
	}

	// feature_detector.cpp:568-583 (cv::goodFeaturesToTrack with the mask of reset())
	Corners detectFeatures(const cv::Mat& image)
	{
		const cv::Mat gray = asGray(image);
		std::vector<double> xy(2 * 8192);
		int n = 0;
		check(ebt_detect_features(ctx_, gray.data, xy.data(), 8192, &n), "ebt_detect_features");
		Corners out;
		for (int i = 0; i < n && i < 8192; ++i) out.emplace_back(xy[2 * i], xy[2 * i + 1]);
		return out;
	}

	// feature_detector.cpp:585-619, buffered
	void updatePatches(const common::EventSample& event)
	{
		ebt_event ev;
		std::memset(&ev, 0, sizeof(ev));
		ev.t_us = event.timestamp.count();
		ev.x = static_cast<uint16_t>(event.value.point.x);
		ev.y = static_cast<uint16_t>(event.value.point.y);
		ev.polarity = static_cast<int8_t>(event.value.sign == common::EventPolarity::POSITIVE ? 1 : -1);
		pending_.push_back(ev);
		if (pending_.size() >= params_.flushThreshold) flush();
	}

	// ---- event motion compensation (GUI only): accepted, not computed ----
	void addEvent(const common::EventSample& event) { lastCompensation_ = event.timestamp; }
	void initMotionField(const common::timestamp_t) {}
	void interpolateMotionField(const common::timestamp_t) {}
	void compensateEvents(const std::list<common::EventSample>&) {}
	void compensateEventsContrast(const std::list<common::EventSample>&) {}
	void integrateEvents(const std::list<common::EventSample>&) {}
	void clearEvents() { lastEvents_.clear(); }
	std::list<common::EventSample> const& getEvents() { return lastEvents_; }
	cv::Mat const& getCompensatedEventImage() { return zeroImage(); }
	cv::Mat const& getIntegratedEventImage() { return zeroImage(); }
	common::timestamp_t const& getLastCompensation() { return lastCompensation_; }

	// feature_detector.cpp:734-742: new parameters, then reset()
	void setParams(const tracker::DetectorParams& params)
	{
		if (ctx_) ebt_destroy(ctx_);
		ctx_ = nullptr;
		params_ = params;
		pending_.clear();
		patches_.clear();
		archivedPatches_.clear();
		corners_.clear();
		create();
	}

	Patches const& getPatches() const
	{
		refresh();
		return patches_;
	}
	Patches& getPatches()
	{
		refresh();
		return patches_;
	}
	Corners const& getFeatures() const { return corners_; }
	Patches const& getArchivedPatches() const
	{
		refresh();
		return archivedPatches_;
	}

	// Optimizer::getFinalCosts (optimizer.h:47): every optimisation, in execution order
	std::vector<tracker::OptimizerFinalLoss> getOptimizedFinalCosts() const
	{
		flush();
		int64_t n = 0;
		check(ebt_get_records(ctx_, nullptr, 0, &n), "ebt_get_records");
		std::vector<ebt_opt_record> r(static_cast<size_t>(n));
		if (n) check(ebt_get_records(ctx_, r.data(), n, &n), "ebt_get_records");
		std::vector<tracker::OptimizerFinalLoss> out;
		for (const ebt_opt_record& k : r) out.push_back({k.track_id, k.final_cost, k.ts});
		return out;
	}

	// pending events to the GPU now (the getters do this themselves)
	void flush() const
	{
		if (pending_.empty()) return;
		check(ebt_push_events(ctx_, pending_.data(), static_cast<int64_t>(pending_.size()), nullptr), "ebt_push_events");
		pending_.clear();
		dirty_ = true;
	}
	ebt_ctx* handle() { return ctx_; }

   private:
	void create()
	{
		if (params_.patchExtent != EBT_PATCH_EXTENT)
			throw std::invalid_argument("tracker::FeatureDetector: patchExtent is fixed at 12 (25 x 25 patches) on the device");
		ebt_params p;
		if (ebt_default_params(&p, params_.imageSize.width, params_.imageSize.height) != EBT_OK)
			throw std::invalid_argument("tracker::FeatureDetector: bad image size");
		p.max_patches = static_cast<int32_t>(params_.maxPatches);
		p.init_num_events = params_.initNumEvents;
		p.association_distance = params_.associationDistance;
		p.max_iterations = params_.optimizerParams.maxNumIterations;
		p.huber_loss = params_.optimizerParams.huberLoss;
		p.optimizer_threshold = params_.optimizerParams.optimizerThreshold;
		if (ebt_create(&p, params_.device, &ctx_) != EBT_OK)
			throw std::runtime_error(std::string("ebt_create: ") + ebt_last_error(nullptr));
		ebt_frontend_params fp;
		check(ebt_default_frontend_params(&fp), "ebt_default_frontend_params");
		fp.quality_level = params_.qualityLevel;
		fp.min_distance = params_.minDistance;
		check(ebt_set_frontend_params(ctx_, &fp), "ebt_set_frontend_params");
		if (params_.optimizerParams.drawCostMap) check(ebt_set_cost_map(ctx_, 1), "ebt_set_cost_map");
		if (params_.drawImages || params_.patchProducts)
			check(ebt_set_products(ctx_, params_.drawImages ? 2 : 1), "ebt_set_products");
		zero_ = cv::Mat::zeros(params_.imageSize.height, params_.imageSize.width, CV_64F);
		dirty_ = true;
	}
	void check(int rc, const char* what) const
	{
		if (rc < 0) throw std::runtime_error(std::string(what) + ": " + ebt_last_error(ctx_));
	}
	cv::Mat const& zeroImage() const { return zero_; }
	cv::Mat asGray(const cv::Mat& image) const
	{
		cv::Mat gray = image;
#ifdef CV_VERSION  // the real OpenCV: accept colour input as the reference's callers may pass it
		if (image.channels() == 3) cv::cvtColor(image, gray, cv::COLOR_BGR2GRAY);
		if (!gray.isContinuous()) gray = gray.clone();
#endif
		if (gray.type() != CV_8U || gray.rows != params_.imageSize.height || gray.cols != params_.imageSize.width ||
			!gray.isContinuous())
			throw std::invalid_argument("tracker::FeatureDetector: images must be CV_8U, continuous, of DetectorParams::imageSize");
		return gray;
	}
	void readFeatures()
	{
		int nc = 0;
		check(ebt_get_features(ctx_, nullptr, 0, &nc), "ebt_get_features");
		std::vector<double> xy(static_cast<size_t>(nc) * 2);
		if (nc) check(ebt_get_features(ctx_, xy.data(), nc, &nc), "ebt_get_features");
		corners_.clear();
		for (int i = 0; i < nc; ++i) corners_.emplace_back(xy[2 * i], xy[2 * i + 1]);
	}
	static cv::Mat toMat(const std::vector<double>& v, int rows, int cols)
	{
		cv::Mat m = cv::Mat::zeros(rows, cols, CV_64F);
		for (int y = 0; y < rows; ++y)
			for (int x = 0; x < cols; ++x) m.at<double>(y, x) = v[static_cast<size_t>(y) * cols + x];
		return m;
	}
	void fill(Patches& dst, int archived) const
	{
		int n = 0;
		check(ebt_get_tracks(ctx_, nullptr, 0, archived, &n), "ebt_get_tracks");
		std::vector<ebt_track> t(static_cast<size_t>(n));
		if (n) check(ebt_get_tracks(ctx_, t.data(), n, archived, &n), "ebt_get_tracks");
		dst.clear();
		const bool products = params_.drawImages || params_.patchProducts;
		std::vector<double> buf(EBT_PATCH_PIXELS);
		for (const ebt_track& k : t)
		{
			Patch p(Corner(k.rect_x + EBT_PATCH_EXTENT, k.rect_y + EBT_PATCH_EXTENT), EBT_PATCH_EXTENT,
					common::timestamp_t(k.current_ts));
			p.v_.has_flow = k.init != 0;
			p.v_.lost = k.lost != 0;
			p.v_.id = k.track_id;
			p.v_.origin = common::Point2d(k.init_x, k.init_y);
			p.v_.last_update = common::timestamp_t(k.time_last_update);
			p.v_.stale_after = common::timestamp_t(k.time_without_update);
			p.v_.created = common::timestamp_t(k.init_time);
			p.v_.window = static_cast<size_t>(k.num_of_events);
			p.v_.flow_angle = k.flow_dir;
			p.v_.warp = common::Pose2d(Sophus::SO2d(k.warp[0], k.warp[1]), Eigen::Vector2d(k.warp[2], k.warp[3]));
			std::vector<double> xyt(static_cast<size_t>(k.n_trajectory) * 3);
			int m = 0;
			check(ebt_get_trajectory(ctx_, k.track_id, xyt.data(), k.n_trajectory, &m), "ebt_get_trajectory");
			p.v_.path.clear();
			for (int i = 0; i < k.n_trajectory; ++i)
				p.v_.path.emplace_back(common::Point2d(xyt[3 * i], xyt[3 * i + 1]),
										   common::timestamp_t(static_cast<int64_t>(xyt[3 * i + 2])));
			p.v_.costs.resize(static_cast<size_t>(k.n_optimizations));
			check(ebt_get_final_costs(ctx_, k.track_id, p.v_.costs.data(), k.n_optimizations, &m), "ebt_get_final_costs");
			check(ebt_get_patch(ctx_, k.track_id, 0, buf.data()), "ebt_get_patch");
			p.v_.events_patch = toMat(buf, EBT_PATCH_SIDE, EBT_PATCH_SIDE);
			if (products)
			{
				check(ebt_get_patch(ctx_, k.track_id, 1, buf.data()), "ebt_get_patch");
				p.v_.compensated_patch = toMat(buf, EBT_PATCH_SIDE, EBT_PATCH_SIDE);
				check(ebt_get_patch(ctx_, k.track_id, 2, buf.data()), "ebt_get_patch");
				p.v_.predicted_patch = toMat(buf, EBT_PATCH_SIDE, EBT_PATCH_SIDE);
			}
			if (params_.optimizerParams.drawCostMap && k.n_optimizations > 0)
			{
				std::vector<double> cm(121);
				check(ebt_get_cost_map(ctx_, k.track_id, cm.data()), "ebt_get_cost_map");
				p.v_.cost_map = toMat(cm, 11, 11);
			}
			dst.push_back(std::move(p));
		}
	}
	void refresh() const
	{
		flush();
		if (!dirty_) return;
		const size_t kept = preExitCopies();
		fill(patches_, 0);
		Patches tail;
		// patches appended by preExit stay behind the device's archive, as in the reference
		auto it = archivedPatches_.end();
		for (size_t i = 0; i < kept; ++i) --it;
		tail.splice(tail.end(), archivedPatches_, it, archivedPatches_.end());
		fill(archivedPatches_, 1);
		archivedPatches_.splice(archivedPatches_.end(), tail);
		archivedFromDevice_ = archivedPatches_.size() - kept;
		dirty_ = false;
	}
	size_t preExitCopies() const { return archivedPatches_.size() - archivedFromDevice_; }

	DetectorParams params_;
	ebt_ctx* ctx_ = nullptr;
	// the reference's getters are const; the flush they imply here is not
	mutable std::vector<ebt_event> pending_;
	mutable Patches patches_;
	mutable Patches archivedPatches_;
	mutable size_t archivedFromDevice_ = 0;
	mutable bool dirty_ = true;
	Corners corners_;
	std::list<common::EventSample> lastEvents_;
	common::timestamp_t lastCompensation_{0};
	cv::Mat zero_;
};
}  // namespace tracker
