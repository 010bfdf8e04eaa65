// TEST INFRASTRUCTURE ONLY — never linked or loaded by the product path.
//
// C-ABI harness around the reference's OWN tracker::Patch and
// tracker::Optimizer (implementation/feature_tracker/src/{patch,optimizer}.cpp,
// compiled unmodified from /root/reference by oracle/Makefile, together with
// the vendored Ceres 2.0 / Sophus / Eigen) so that tests and bench.py's
// cpu_baseline leg can drive the real reference code.  The per-event driver
// (FeatureDetector::updatePatches / updateNumOfEvents / newImage /
// associatePatches) is restated here line by line from
// implementation/feature_tracker/src/feature_detector.cpp:493-732 because that
// file needs OpenCV's detector/LK/Sobel and the motion-compensation code;
// corners and LK flows are injected by the caller instead.
//
// Output: oracle/_ref/libebt_ref.so (git-ignored, travels to the GPU box).

#include <feature_tracker/local_parameterization_se2.hpp>
#include <feature_tracker/optimizer.h>
#include <feature_tracker/patch.h>
#include <opencv2/core/eigen.hpp>

#include <spdlog/sinks/null_sink.h>

#include <atomic>
#include <chrono>
#include <map>
#include <thread>
#include <unordered_map>

#include "../include/ebt_types.h"

namespace
{
void ensureLoggers()
{
	// every reference class does spdlog::get("console") and dereferences it
	if (!spdlog::get("console"))
	{
		auto sink = std::make_shared<spdlog::sinks::null_sink_mt>();
		auto a = std::make_shared<spdlog::logger>("console", sink);
		auto b = std::make_shared<spdlog::logger>("stderr", sink);
		a->set_level(spdlog::level::off);
		b->set_level(spdlog::level::off);
		spdlog::register_logger(a);
		spdlog::register_logger(b);
	}
}

common::EventSample toSample(const ebt_event& e)
{
	common::EventSample s;
	s.value.point = common::Point2i(e.x, e.y);
	s.value.sign = e.polarity > 0 ? common::POSITIVE : common::NEGATIVE;
	s.timestamp = common::timestamp_t(e.t_us);
	return s;
}

// FeatureDetector::getLogImage + getGradients (feature_detector.cpp:714-732):
// L = log(I * (1/255) + 0.1); Sobel 3x3 of L/8, BORDER_REFLECT_101.
// Separable order as OpenCV's: row pass first, then column pass.
void logSobel(const uint8_t* gray, int w, int h, cv::Mat& gx, cv::Mat& gy)
{
	double lut[256];
	for (int i = 0; i < 256; ++i)
		lut[i] = std::log(i * (1.0 / 255.0) + 10e-2) * (1.0 / 8);
	std::vector<double> L(static_cast<size_t>(w) * h);
	for (size_t i = 0; i < L.size(); ++i) L[i] = lut[gray[i]];
	auto refl = [](int i, int n) {
		if (i < 0) return -i;
		if (i >= n) return 2 * n - 2 - i;
		return i;
	};
	std::vector<double> dif(L.size()), smo(L.size());
	for (int y = 0; y < h; ++y)
		for (int x = 0; x < w; ++x)
		{
			const double a = L[y * w + refl(x - 1, w)];
			const double b = L[y * w + x];
			const double c = L[y * w + refl(x + 1, w)];
			dif[y * w + x] = c - a;
			smo[y * w + x] = a + b * 2 + c;
		}
	gx = cv::Mat(h, w, CV_64F);
	gy = cv::Mat(h, w, CV_64F);
	for (int y = 0; y < h; ++y)
		for (int x = 0; x < w; ++x)
		{
			const int y0 = refl(y - 1, h), y2 = refl(y + 1, h);
			gx.at<double>(y, x) =
				dif[y0 * w + x] + dif[y * w + x] * 2 + dif[y2 * w + x];
			gy.at<double>(y, x) = smo[y2 * w + x] - smo[y0 * w + x];
		}
}

// OptimizerParams::drawCostMap (optimizer.h:18) for every Optimizer created
// after the call: Optimizer::drawCostMap (optimizer.cpp:32-60) then runs inside
// optimize() and Patch::getCostMap() returns the 11x11 map.
bool g_drawCostMap = false;
double* g_costMapsOut = nullptr;  // ref_optimize_batch_cm: n x 121 doubles

struct RefTracker
{
	// mirrors of FeatureDetector members (feature_detector.h:94-118)
	// the DetectorParams fields this path reads, with the reference's
	// defaults (feature_detector.h:10-31); feature_detector.h itself pulls
	// in the OpenCV LK/detector headers, so it is not included.
	struct
	{
		double associationDistance = 5;
		int32_t patchExtent = 12;
		cv::Size imageSize = {240, 180};
		tracker::OptimizerParams optimizerParams = {};
		int initNumEvents = 75;
		size_t maxPatches = 100;
	} params;
	cv::Mat gradX, gradY;  // latest frame
	std::unordered_map<size_t, std::unique_ptr<tracker::Optimizer>> optimizers;
	tracker::Patches patches;
	tracker::Patches archived;
	size_t nextTrackId = 0;
	int imageCounter = 0;  // FlowEstimator::imageCounter_
	bool callWarpImage = false;
	bool drawImages = false;  // DetectorParams::drawImages (feature_detector.cpp:504-510)
	std::vector<ebt_opt_record> records;
	int64_t eventIndex = 0;
};

// FeatureDetector::updateNumOfEvents (feature_detector.cpp:666-712).  The
// reference warps both full gradient frames and then crops the rect; only the
// cropped 25x25 texels are used, so this evaluates cv::warpAffine's nearest
// rule (ref_shim/opencv2/opencv.hpp) on the crop alone — same values.
void updateNumOfEvents(RefTracker& T, tracker::Patch& patch)
{
	const auto rect = patch.getPatch();
	const auto center = (rect.tl() + rect.br()) * 0.5;
	if (center.x <= 5 || center.y <= 5 || center.x >= T.gradX.cols - 5 ||
		center.y >= T.gradX.rows - 5)
	{
		patch.setLost();
		return;
	}
	if (rect.x < 0 || rect.y < 0 || rect.x + rect.width >= T.gradX.cols ||
		rect.y + rect.height >= T.gradX.rows)
	{
		patch.setNumOfEvents(T.params.initNumEvents);
		return;
	}
	const auto M = patch.getWarp().matrix2x3();
	const cv::Rect r(rect);  // rounds x, y, width, height
	const auto flow = patch.getFlow();  // float
	const double a = 0.6 * std::cos(flow);  // std::cos(float) -> float
	const double b = 0.6 * std::sin(flow);
	double acc = 0;
	for (int y = r.y; y < r.y + r.height; ++y)
		for (int x = r.x; x < r.x + r.width; ++x)
		{
			const int X0 = cv::shimRound(M(0, 0) * x * 1024) +
						   cv::shimRound((M(0, 1) * y + M(0, 2)) * 1024) + 512;
			const int Y0 = cv::shimRound(M(1, 0) * x * 1024) +
						   cv::shimRound((M(1, 1) * y + M(1, 2)) * 1024) + 512;
			const int xi = X0 >> 10, yi = Y0 >> 10;
			double vx = 0, vy = 0;
			if (xi >= 0 && xi < T.gradX.cols && yi >= 0 && yi < T.gradX.rows)
			{
				vx = T.gradX.at<double>(yi, xi);
				vy = T.gradY.at<double>(yi, xi);
			}
			acc += std::fabs(vx * a + vy * b);
		}
	const size_t sumPatch = acc;
	patch.setNumOfEvents(sumPatch);
}

void fillTrack(const tracker::Patch& cp, ebt_track& t)
{
	auto& p = const_cast<tracker::Patch&>(cp);  // isInit() is non-const
	std::memset(&t, 0, sizeof(t));
	t.track_id = p.getTrackId();
	t.lost = p.isLost();
	t.init = p.isInit();
	t.num_of_events = static_cast<int32_t>(p.getNumOfEvents());
	t.window_size = static_cast<int32_t>(p.getEvents().size());
	t.n_optimizations = static_cast<int32_t>(p.getFinalCosts().size());
	t.n_trajectory = static_cast<int32_t>(p.getTrajectory().size());
	t.rect_x = p.getPatch().x;
	t.rect_y = p.getPatch().y;
	const auto ip = p.getInitPatch();
	t.init_x = ip.x + (ip.width - 1) / 2.;
	t.init_y = ip.y + (ip.height - 1) / 2.;
	if (p.isInit())
	{
		std::memcpy(t.warp, p.getWarp().data(), 4 * sizeof(double));
		t.flow_dir = p.getFlow();
	}
	else
	{
		t.warp[0] = 1;
	}
	t.last_cost = p.getFinalCosts().empty() ? 0 : p.getFinalCosts().back();
	t.current_ts = p.getCurrentTimestamp().count();
	t.time_last_update = t.n_optimizations ? p.getTimeLastUpdate().count() : 0;
	t.time_without_update = p.getTimeWithoutUpdate().count();
	t.init_time = p.getInitTime().count();
	t.counter = -1;  // Patch::counter_ has no getter
}
}  // namespace

extern "C" {

void* ref_create(int width, int height, int maxPatches, int callWarpImage)
{
	ensureLoggers();
	auto* T = new RefTracker();
	T->params.imageSize = cv::Size(width, height);
	T->params.maxPatches = static_cast<size_t>(maxPatches);
	T->callWarpImage = callWarpImage != 0;
	return T;
}

void ref_destroy(void* h) { delete static_cast<RefTracker*>(h); }
void ref_set_draw_images(void* h, int on) { static_cast<RefTracker*>(h)->drawImages = on != 0; }

// FeatureDetector::newImage part 1 = extractPatches (feature_detector.cpp:
// 543-566) with the Harris corners injected, then associatePatches (:630-664).
void ref_new_frame_begin(void* h, const uint8_t* gray, int64_t ts,
						 const double* cornersXY, int nCorners)
{
	auto& T = *static_cast<RefTracker*>(h);
	const auto stamp = common::timestamp_t(ts);
	tracker::Patches newPatches;
	for (int i = 0; i < nCorners; ++i)
	{
		newPatches.emplace_back(tracker::Patch(
			tracker::Corner(cornersXY[2 * i], cornersXY[2 * i + 1]),
			T.params.patchExtent, stamp));
	}
	logSobel(gray, T.params.imageSize.width, T.params.imageSize.height, T.gradX,
			 T.gradY);
	T.params.optimizerParams.drawCostMap = g_drawCostMap;
	T.optimizers[ts] = std::make_unique<tracker::Optimizer>(
		tracker::Optimizer(T.params.optimizerParams, T.params.imageSize));
	T.optimizers[ts]->setGrad(T.gradX, T.gradY);

	// associatePatches
	for (auto& patch : T.patches)
	{
		const auto& corner = patch.toCorner();
		for (auto& newPatch : newPatches)
		{
			const auto& newCorner = newPatch.toCorner();
			if (newPatch.getTrackId() == -1 &&
				cv::norm(corner - newCorner) < T.params.associationDistance)
			{
				newPatch.setTrackId(patch.getTrackId());
				break;
			}
		}
		patch.setTs(stamp);
		patch.addTrajectoryPosition();
	}
	for (auto& newPatch : newPatches)
	{
		if (newPatch.getTrackId() == -1 &&
			T.patches.size() < T.params.maxPatches)
		{
			newPatch.setTrackId(T.nextTrackId);
			newPatch.setGrad(T.gradX, T.gradY);
			T.patches.push_back(newPatch);
			T.optimizers[ts]->addUser();
			T.nextTrackId++;
		}
	}
	if (T.imageCounter < 2) T.imageCounter++;  // FlowEstimator::addImage
}

// Patches FlowEstimator::getFlowPatches would query LK for
// (flow_estimator.cpp:27-39): returns 0 before the second image.
int ref_get_uninit(void* h, int32_t* ids, double* cornersXY, int cap)
{
	auto& T = *static_cast<RefTracker*>(h);
	if (T.imageCounter != 2) return 0;
	int n = 0;
	for (auto& patch : T.patches)
	{
		if (patch.isInit()) continue;
		if (n < cap)
		{
			ids[n] = patch.getTrackId();
			cornersXY[2 * n] = patch.toCorner().x;
			cornersXY[2 * n + 1] = patch.toCorner().y;
		}
		++n;
	}
	return n;
}

// Loop body of FlowEstimator::getFlowPatches (flow_estimator.cpp:41-79) with
// the LK result (status, cv::Point2f nextPoint) injected.
void ref_set_flow(void* h, int32_t trackId, int found, float nextX, float nextY)
{
	auto& T = *static_cast<RefTracker*>(h);
	for (auto& patch : T.patches)
	{
		if (patch.getTrackId() != trackId || patch.isInit()) continue;
		if (!found)
		{
			patch.setLost();
			return;
		}
		const auto& corner = patch.toCorner();
		const auto dirX = nextX - corner.x;
		const auto dirY = nextY - corner.y;
		const auto flowDir = std::atan2(dirY, dirX);
		common::Pose2d warp;
		warp.translation() = Eigen::Vector2d(-dirX, -dirY);
		patch.setWarp(warp);
		patch.setFlowDir(flowDir);
		cv::Point2d flow(dirX, dirY);
		patch.setTimeWithoutUpdate(common::timestamp_t(
			static_cast<int64_t>(1e6 / fmax(1e-1, cv::norm(flow)))));
		const auto newCorner = patch.toCorner();
		if (newCorner.x <= 5 || newCorner.y <= 5 ||
			newCorner.x >= T.params.imageSize.width - 5 ||
			newCorner.y >= T.params.imageSize.height - 5)
		{
			patch.setLost();
		}
		return;
	}
}

// FeatureDetector::newImage part 2 (feature_detector.cpp:499-535).
void ref_new_frame_end(void* h)
{
	auto& T = *static_cast<RefTracker*>(h);
	for (auto& patch : T.patches) updateNumOfEvents(T, patch);
	if (T.drawImages)
		for (auto& patch : T.patches) patch.warpImage();
	for (auto it = T.patches.begin(); it != T.patches.end();)
	{
		if (it->isLost())
		{
			T.archived.push_back(*it);
			T.optimizers[it->getInitTime().count()]->deleteUser();
			it = T.patches.erase(it);
			continue;
		}
		++it;
	}
	for (auto it = T.optimizers.begin(); it != T.optimizers.end();)
	{
		if (!it->second->isUsed())
			it = T.optimizers.erase(it);
		else
			++it;
	}
}

// FeatureDetector::updatePatches (feature_detector.cpp:585-619), one call per
// event in stream order.  Returns the number of Optimizer::optimize calls.
int64_t ref_push_events(void* h, const ebt_event* ev, int64_t n)
{
	auto& T = *static_cast<RefTracker*>(h);
	int64_t nOpt = 0;
	for (int64_t i = 0; i < n; ++i, ++T.eventIndex)
	{
		const auto event = toSample(ev[i]);
		for (auto& patch : T.patches)
		{
			if (patch.isLost()) continue;
			if (patch.isInPatch(event.value.point)) patch.addEvent(event);
			if (patch.isReady() && patch.isInit())
			{
				const size_t nBefore = patch.getFinalCosts().size();
				T.optimizers[patch.getInitTime().count()]->optimize(patch);
				updateNumOfEvents(T, patch);
				if (T.callWarpImage) patch.warpImage();
				++nOpt;
				ebt_opt_record r;
				std::memset(&r, 0, sizeof(r));
				r.track_id = patch.getTrackId();
				r.n_successful = r.n_unsuccessful = r.termination = -1;
				r.lost = patch.isLost();
				r.num_of_events = static_cast<int32_t>(patch.getNumOfEvents());
				r.ts = patch.getCurrentTimestamp().count();
				r.event_index = T.eventIndex;
				r.x = patch.toCorner().x;
				r.y = patch.toCorner().y;
				r.final_cost = patch.getFinalCosts().size() > nBefore
								   ? patch.getFinalCosts().back()
								   : 0;
				std::memcpy(r.warp, patch.getWarp().data(), 4 * sizeof(double));
				r.flow_dir = patch.getFlow();
				T.records.push_back(r);
			}
		}
	}
	return nOpt;
}

int ref_num_patches(void* h)
{
	return static_cast<int>(static_cast<RefTracker*>(h)->patches.size());
}
int ref_num_archived(void* h)
{
	return static_cast<int>(static_cast<RefTracker*>(h)->archived.size());
}
int ref_get_tracks(void* h, ebt_track* out, int cap, int archived)
{
	auto& T = *static_cast<RefTracker*>(h);
	int n = 0;
	for (const auto& p : archived ? T.archived : T.patches)
	{
		if (n < cap) fillTrack(p, out[n]);
		++n;
	}
	return n;
}
int64_t ref_num_records(void* h)
{
	return static_cast<int64_t>(static_cast<RefTracker*>(h)->records.size());
}
int64_t ref_get_records(void* h, ebt_opt_record* out, int64_t cap)
{
	auto& T = *static_cast<RefTracker*>(h);
	const int64_t n = std::min<int64_t>(cap, T.records.size());
	std::memcpy(out, T.records.data(), n * sizeof(ebt_opt_record));
	return n;
}
void ref_clear_records(void* h) { static_cast<RefTracker*>(h)->records.clear(); }

static tracker::Patch* findPatch(RefTracker& T, int32_t id)
{
	for (auto& p : T.patches)
		if (p.getTrackId() == id) return &p;
	for (auto& p : T.archived)
		if (p.getTrackId() == id) return &p;
	return nullptr;
}
// which: 0 integratedNabla, 1 motion-compensated, 2 predictedNabla, 3 costMap
int ref_get_patch(void* h, int32_t id, int which, double* out625)
{
	auto* p = findPatch(*static_cast<RefTracker*>(h), id);
	if (!p) return -1;
	const cv::Mat& m = which == 0   ? p->getIntegratedNabla()
					   : which == 1 ? p->getCompenatedIntegratedNabla()
					   : which == 2 ? p->getPredictedNabla()
									: p->getCostMap();
	for (int y = 0; y < m.rows; ++y)
		for (int x = 0; x < m.cols; ++x) out625[y * m.cols + x] = m.at<double>(y, x);
	return m.rows * m.cols;
}
// trajectory as (x, y, ts) triples of doubles
int ref_get_trajectory(void* h, int32_t id, double* out, int cap)
{
	auto* p = findPatch(*static_cast<RefTracker*>(h), id);
	if (!p) return -1;
	int n = 0;
	for (const auto& s : p->getTrajectory())
	{
		if (n < cap)
		{
			out[3 * n] = s.value.x;
			out[3 * n + 1] = s.value.y;
			out[3 * n + 2] = static_cast<double>(s.timestamp.count());
		}
		++n;
	}
	return n;
}
int ref_get_gradients(void* h, double* gx, double* gy)
{
	auto& T = *static_cast<RefTracker*>(h);
	for (int y = 0; y < T.gradX.rows; ++y)
		for (int x = 0; x < T.gradX.cols; ++x)
		{
			gx[y * T.gradX.cols + x] = T.gradX.at<double>(y, x);
			gy[y * T.gradX.cols + x] = T.gradY.at<double>(y, x);
		}
	return T.gradX.rows * T.gradX.cols;
}
}  // extern "C"

// ---------------------------------------------------------------------------
// Batched entry: F independent feature-patch problems, each one real
// Patch + real Optimizer::optimize (optimizer.cpp:62-206).  Used for kernel
// level parity and as the CPU baseline ("reference" kind).
// ---------------------------------------------------------------------------
namespace
{
struct RefFrames
{
	int width = 0, height = 0;
	std::vector<cv::Mat> gx, gy;
};

// Build a Patch whose state equals (initPoint, warp, flow, window): the rect
// is what Patch::updatePatchRect (patch.cpp:49-63) derives from the warp.
tracker::Patch makePatch(const ebt_problem& pr, const ebt_event* events)
{
	tracker::Patch patch(tracker::Corner(pr.init_x, pr.init_y), EBT_PATCH_EXTENT,
						 common::timestamp_t(0));
	common::Pose2d warp;
	std::memcpy(warp.data(), pr.warp, 4 * sizeof(double));
	patch.setWarp(warp);
	patch.setFlowDir(pr.flow_dir);
	patch.updatePatchRect();
	patch.setNumOfEvents(static_cast<size_t>(pr.n_events));
	// newest first in the array == deque order; addEvent pushes to the front
	for (int i = pr.n_events - 1; i >= 0; --i)
		patch.addEvent(toSample(events[pr.event_offset + i]));
	return patch;
}

void solveRange(const RefFrames* F, const ebt_problem* problems,
				const ebt_event* events, ebt_solution* out, double* nabla,
				std::atomic<int64_t>* next, int64_t n)
{
	std::vector<std::unique_ptr<tracker::Optimizer>> opts(F->gx.size());
	for (;;)
	{
		const int64_t i = next->fetch_add(1);
		if (i >= n) break;
		const ebt_problem& pr = problems[i];
		auto& opt = opts[pr.frame_slot];
		if (!opt)
		{
			tracker::OptimizerParams op;
			op.drawCostMap = g_drawCostMap;
			opt.reset(new tracker::Optimizer(op, cv::Size(F->width, F->height)));
			opt->setGrad(F->gx[pr.frame_slot], F->gy[pr.frame_slot]);
		}
		tracker::Patch patch = makePatch(pr, events);
		opt->optimize(patch);
		ebt_solution& s = out[i];
		std::memset(&s, 0, sizeof(s));
		std::memcpy(s.warp, patch.getWarp().data(), 4 * sizeof(double));
		s.flow_dir = patch.getFlow();
		s.final_cost = patch.getFinalCosts().back();
		s.x = patch.toCorner().x;
		s.y = patch.toCorner().y;
		s.n_successful = s.n_unsuccessful = s.termination = -1;
		s.ts = patch.getCurrentTimestamp().count();
		const cv::Mat& m = patch.getIntegratedNabla();
		int inRect = 0;
		for (int y = 0; y < EBT_PATCH_SIDE; ++y)
			for (int x = 0; x < EBT_PATCH_SIDE; ++x)
			{
				const double v = m.at<double>(y, x);
				if (nabla) nabla[i * EBT_PATCH_PIXELS + y * EBT_PATCH_SIDE + x] = v;
				(void)inRect;
			}
		s.n_passes = -1;
		if (g_costMapsOut)
		{
			// zeros(1, 1) until drawCostMap ran (patch.cpp:29)
			const cv::Mat& cm = patch.getCostMap();
			for (int k = 0; k < 121; ++k)
				g_costMapsOut[i * 121 + k] =
					(cm.rows == 11 && cm.cols == 11) ? cm.at<double>(k / 11, k % 11) : 0.0;
		}
	}
}
}  // namespace

extern "C" {
void* ref_frames_create(int width, int height)
{
	ensureLoggers();
	auto* F = new RefFrames();
	F->width = width;
	F->height = height;
	return F;
}
void ref_frames_destroy(void* f) { delete static_cast<RefFrames*>(f); }
// returns the slot index of the new template frame
int ref_frames_add(void* f, const uint8_t* gray)
{
	auto& F = *static_cast<RefFrames*>(f);
	cv::Mat gx, gy;
	logSobel(gray, F.width, F.height, gx, gy);
	F.gx.push_back(gx);
	F.gy.push_back(gy);
	return static_cast<int>(F.gx.size()) - 1;
}
int ref_frames_get(void* f, int slot, double* gx, double* gy)
{
	auto& F = *static_cast<RefFrames*>(f);
	for (int y = 0; y < F.height; ++y)
		for (int x = 0; x < F.width; ++x)
		{
			gx[y * F.width + x] = F.gx[slot].at<double>(y, x);
			gy[y * F.width + x] = F.gy[slot].at<double>(y, x);
		}
	return 0;
}

// nabla (optional): n x 625 doubles, the integer event patches.
// Returns wall seconds spent inside the solves.
double ref_optimize_batch(void* f, const ebt_problem* problems, int64_t n,
						  const ebt_event* events, ebt_solution* out,
						  double* nabla, int threads)
{
	auto* F = static_cast<RefFrames*>(f);
	std::atomic<int64_t> next(0);
	const auto t0 = std::chrono::steady_clock::now();
	if (threads <= 1)
	{
		solveRange(F, problems, events, out, nabla, &next, n);
	}
	else
	{
		std::vector<std::thread> pool;
		for (int t = 0; t < threads; ++t)
			pool.emplace_back(solveRange, F, problems, events, out, nabla, &next,
							  n);
		for (auto& th : pool) th.join();
	}
	return std::chrono::duration<double>(std::chrono::steady_clock::now() - t0)
		.count();
}

void ref_set_draw_cost_map(int on) { g_drawCostMap = on != 0; }
// ref_optimize_batch with OptimizerParams::drawCostMap on: cost_maps = n x 121
// doubles, row-major costMap.at<double>(y + 5, x + 5).
double ref_optimize_batch_cm(void* f, const ebt_problem* problems, int64_t n,
							 const ebt_event* events, ebt_solution* out,
							 double* nabla, double* cost_maps)
{
	const bool was = g_drawCostMap;
	g_drawCostMap = true;
	g_costMapsOut = cost_maps;
	const double t = ref_optimize_batch(f, problems, n, events, out, nabla, 1);
	g_costMapsOut = nullptr;
	g_drawCostMap = was;
	return t;
}

// Residuals (625) and Ceres-autodiff Jacobians of the reference functor
// (optimizer_cost.h:29-90) at (warp, flow): jWarp 625x4 row-major (ambient
// cos,sin,tx,ty), jFlow 625, jLocal 625x3 = jWarp * Dx_this_mul_exp_x_at_0.
// m625 = normalised integrated patch; rect given by its top-left corner.
int ref_eval_functor(void* f, int slot, const double* m625, double rectX,
					 double rectY, const double* warp4, double flow,
					 double* residuals, double* jWarp, double* jFlow,
					 double* jLocal)
{
	auto& F = *static_cast<RefFrames*>(f);
	std::vector<double> grad(static_cast<size_t>(F.width) * F.height * 2);
	for (int y = 0; y < F.height; ++y)
		for (int x = 0; x < F.width; ++x)
		{
			grad[2 * (y * F.width + x)] = F.gx[slot].at<double>(y, x);
			grad[2 * (y * F.width + x) + 1] = F.gy[slot].at<double>(y, x);
		}
	tracker::Grid grid(grad.data(), 0, F.height, 0, F.width);
	tracker::Interpolator interp(grid);
	cv::Mat m(EBT_PATCH_SIDE, EBT_PATCH_SIDE, CV_64F);
	std::memcpy(m.data, m625, EBT_PATCH_PIXELS * sizeof(double));
	auto* c = new tracker::OptimizerCostFunctor(
		m, &interp, cv::Rect2d(rectX, rectY, EBT_PATCH_SIDE, EBT_PATCH_SIDE),
		cv::Size(F.width, F.height));
	ceres::AutoDiffCostFunction<tracker::OptimizerCostFunctor, ceres::DYNAMIC, 4,
								1>
		cost(c, EBT_PATCH_PIXELS);
	const double* params[2] = {warp4, &flow};
	double* jac[2] = {jWarp, jFlow};
	if (!cost.Evaluate(params, residuals, jac)) return -1;
	if (jLocal)
	{
		Sophus::SE2d T;
		std::memcpy(T.data(), warp4, 4 * sizeof(double));
		const Eigen::Matrix<double, 4, 3> D = T.Dx_this_mul_exp_x_at_0();
		for (int i = 0; i < EBT_PATCH_PIXELS; ++i)
			for (int k = 0; k < 3; ++k)
			{
				double acc = 0;
				for (int a = 0; a < 4; ++a) acc += jWarp[i * 4 + a] * D(a, k);
				jLocal[i * 3 + k] = acc;
			}
	}
	return 0;
}

// cv::warpAffine nearest index map of the shim for a 2x3 dst->src matrix:
// used to pin the fixed-point rule against cv2 (tests/golden).
void ref_warp_index_map(const double* M6, int w, int h, int32_t* xi, int32_t* yi)
{
	for (int y = 0; y < h; ++y)
		for (int x = 0; x < w; ++x)
		{
			const int X0 = cv::shimRound(M6[0] * x * 1024) +
						   cv::shimRound((M6[1] * y + M6[2]) * 1024) + 512;
			const int Y0 = cv::shimRound(M6[3] * x * 1024) +
						   cv::shimRound((M6[4] * y + M6[5]) * 1024) + 512;
			xi[y * w + x] = X0 >> 10;
			yi[y * w + x] = Y0 >> 10;
		}
}
}  // extern "C"

// Replays the body of the reference's own known-answer test
// Patch.integrateEventsTest (implementation/feature_tracker/test/
// patch_test.cpp:35-60) for an arbitrary corner/extent/event list: construct,
// addEvent each, integrateEvents, return the integrated patch.
extern "C" int ref_patch_integrate(double cx, double cy, int extent,
								   const ebt_event* ev, int n, double* out,
								   int64_t* frontBackTs)
{
	ensureLoggers();
	tracker::Patch patch(tracker::Corner(cx, cy), extent, common::timestamp_t(0));
	for (int i = 0; i < n; ++i) patch.addEvent(toSample(ev[i]));
	patch.integrateEvents();
	const cv::Mat& m = patch.getIntegratedNabla();
	for (int y = 0; y < m.rows; ++y)
		for (int x = 0; x < m.cols; ++x) out[y * m.cols + x] = m.at<double>(y, x);
	if (frontBackTs)
	{
		frontBackTs[0] = patch.getEvents().front().timestamp.count();
		frontBackTs[1] = patch.getEvents().back().timestamp.count();
		frontBackTs[2] = patch.getCurrentTimestamp().count();
		frontBackTs[3] = patch.isReady();
	}
	return m.rows * m.cols;
}

// common::CameraModel<double> (common/include/common/camera_model.h), the
// reference's own header: unproject / project for n points.
#include <common/camera_model.h>
extern "C" void ref_unproject(const ebt_camera* cam, const double* xy, int n, double* out3)
{
	common::CameraModelParams<double> p;
	p.fx = cam->fx; p.fy = cam->fy; p.cx = cam->cx; p.cy = cam->cy;
	p.k1 = cam->k1; p.k2 = cam->k2; p.k3 = cam->k3; p.p1 = cam->p1; p.p2 = cam->p2;
	common::CameraModel<double> m(p);
	for (int i = 0; i < n; ++i)
	{
		const Eigen::Vector3d b = m.unproject(Eigen::Vector2d(xy[2 * i], xy[2 * i + 1]));
		out3[3 * i] = b[0]; out3[3 * i + 1] = b[1]; out3[3 * i + 2] = b[2];
	}
}
extern "C" void ref_project(const ebt_camera* cam, const double* xyz, int n, double* out2)
{
	common::CameraModelParams<double> p;
	p.fx = cam->fx; p.fy = cam->fy; p.cx = cam->cx; p.cy = cam->cy;
	p.k1 = cam->k1; p.k2 = cam->k2; p.k3 = cam->k3; p.p1 = cam->p1; p.p2 = cam->p2;
	common::CameraModel<double> m(p);
	for (int i = 0; i < n; ++i)
	{
		const Eigen::Vector2d r = m.project(Eigen::Vector3d(xyz[3 * i], xyz[3 * i + 1], xyz[3 * i + 2]));
		out2[2 * i] = r[0]; out2[2 * i + 1] = r[1];
	}
}
