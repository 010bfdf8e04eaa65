// TEST INFRASTRUCTURE ONLY (oracle/): the CPU side of BASELINE.json configs[4] — the reference's
// visual-odometry back end (implementation/visual_odometry/src/*.cpp, vendored OpenGV and Ceres),
// compiled UNMODIFIED against the drop-in tracker headers (include/ebt_dropin) and driven the way
// tools::Evaluator drives it:
//
//   eventCallback  tools/evaluator/src/evaluator.cpp:32-44   addEvent, updatePatches, the compensation calls
//   imageCallback  tools/evaluator/src/evaluator.cpp:50-104  newImage, getFeatures; from the third image on
//                  Keyframe(getPatches(), timestamp) -> VisualOdometryFrontEnd::newKeyframeCandidate
//   ~Evaluator     evaluator.cpp:15-21                        preExit, getArchivedPatches, getOptimizedFinalCosts
//
// Two modes over one binary input file (written by tests/test_c5_vo.py):
//   gpu     frames + events  -> tracker::FeatureDetector (drop-in: CUDA library behind the reference API)
//   tracks  per-image (track id, corner) lists of ANY tracker (the CPU oracle's, for the comparison) ->
//           tracker::Patch snapshots -> the same Keyframe / VisualOdometryFrontEnd path
// Output (text): keyframes stored + active, the ATE of visual_odometry::align_cameras_sim3
// (vo/src/aligner.cpp:95-114) against the ground-truth poses, and the camera centres.
#include <feature_tracker/feature_detector.h>
#include <visual_odometry/aligner.h>
#include <visual_odometry/visual_odometry.h>

#include <spdlog/sinks/null_sink.h>

#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <memory>
#include <string>
#include <vector>

namespace vo = visual_odometry;

// OpenGV seeds every RANSAC with time(0) + clock() (opengv/sac/implementation/SampleConsensusProblem.hpp:
// 43-44; the reference leaves randomSeed at its default), so the reference's back end differs from run to
// run on the SAME tracks.  The sources stay unmodified; this executable pins the two C clocks instead
// (its own definitions take precedence over libc's), so that two runs on equal tracks draw equal samples.
#include <ctime>
extern "C" time_t time(time_t* t)
{
	if (t) *t = 1;
	return 1;
}
extern "C" clock_t clock(void) { return 1; }

template <typename T>
static std::vector<T> readVec(FILE* f)
{
	int64_t n = 0;
	if (std::fread(&n, sizeof(n), 1, f) != 1) std::exit(3);
	std::vector<T> v(static_cast<size_t>(n));
	if (n && std::fread(v.data(), sizeof(T), v.size(), f) != v.size()) std::exit(3);
	return v;
}

int main(int argc, char** argv)
{
	if (argc < 4) return 2;
	const std::string mode = argv[1];
	FILE* f = std::fopen(argv[2], "rb");
	if (!f) return 2;
	// the reference's classes fetch these two loggers by name (ft/test/optimizer_test.cpp:10-20)
	auto sink = std::make_shared<spdlog::sinks::null_sink_mt>();
	spdlog::register_logger(std::make_shared<spdlog::logger>("console", sink));
	spdlog::register_logger(std::make_shared<spdlog::logger>("stderr", sink));

	const auto hdr = readVec<int64_t>(f);  // width, height, nFrames, maxPatches
	const int W = static_cast<int>(hdr[0]), H = static_cast<int>(hdr[1]), nF = static_cast<int>(hdr[2]);
	const auto calib = readVec<double>(f);    // fx fy cx cy k1 k2 k3 p1 p2
	const auto frameTs = readVec<int64_t>(f);
	const auto gtRows = readVec<double>(f);   // rows: t_us tx ty tz qx qy qz qw
	common::CameraModelParams<double> cam;
	cam.fx = calib[0]; cam.fy = calib[1]; cam.cx = calib[2]; cam.cy = calib[3];
	cam.k1 = calib[4]; cam.k2 = calib[5]; cam.k3 = calib[6]; cam.p1 = calib[7]; cam.p2 = calib[8];
	common::GroundTruth gt;
	for (size_t r = 0; r + 7 < gtRows.size() + 1 && r + 8 <= gtRows.size(); r += 8)
	{
		const Eigen::Quaterniond q(gtRows[r + 7], gtRows[r + 4], gtRows[r + 5], gtRows[r + 6]);
		gt.emplace_back(common::Pose3d(q.normalized(), Eigen::Vector3d(gtRows[r + 1], gtRows[r + 2], gtRows[r + 3])),
						common::timestamp_t(static_cast<int64_t>(gtRows[r])));
	}
	vo::VisualOdometryFrontEnd visualOdometry(cam, vo::VisualOdometryParams());
	visualOdometry.setGroundTruthSamples(gt);
	// optional: the live patches after every image (image index, track id, corner), for the test's comparison
	FILE* dump = argc > 4 ? std::fopen(argv[4], "w") : nullptr;
	size_t imageNum = 0;
	size_t nTracksLast = 0, nOptimisations = 0, nArchived = 0;
	std::srand(1);  // OpenGV's RANSAC draws from rand()

	if (mode == "gpu")
	{
		const auto frames = readVec<uint8_t>(f);
		const auto events = readVec<ebt_event>(f);
		std::fclose(f);
		tracker::DetectorParams params;  // Evaluator::reset (evaluator.cpp:104-118)
		params.drawImages = false;
		params.imageSize = {W, H};
		params.maxPatches = static_cast<size_t>(hdr[3]);
		std::unique_ptr<tracker::FeatureDetector> tracker_(new tracker::FeatureDetector(params));
		tracker::Corners corners_;
		size_t e = 0;
		for (int k = 0; k < nF; ++k)
		{
			// ---- Replayer::next (replayer.cpp:56-98): events older than the image first
			for (; e < events.size() && events[e].t_us < frameTs[k]; ++e)
			{
				common::EventSample sample;
				sample.timestamp = common::timestamp_t(events[e].t_us);
				sample.value.point = {events[e].x, events[e].y};
				sample.value.sign = events[e].polarity > 0 ? common::EventPolarity::POSITIVE : common::EventPolarity::NEGATIVE;
				// ---- Evaluator::eventCallback (evaluator.cpp:32-44)
				tracker_->addEvent(sample);
				tracker_->updatePatches(sample);
				if ((sample.timestamp - tracker_->getLastCompensation()).count() >= 1000000000 or
					tracker_->getEvents().size() >= 15000)
				{
					tracker_->compensateEventsContrast(tracker_->getEvents());
					tracker_->integrateEvents(tracker_->getEvents());
					tracker_->clearEvents();
				}
			}
			// ---- Evaluator::imageCallback (evaluator.cpp:50-104), tracker + VO branch
			unsigned char* px = const_cast<unsigned char*>(frames.data()) + static_cast<size_t>(k) * W * H;
			common::ImageSample sample(cv::Mat(H, W, CV_8U, px), common::timestamp_t(frameTs[k]));
			imageNum++;
			tracker_->newImage(sample);
			corners_ = tracker_->getFeatures();
			if (imageNum > 2)
			{
				auto keyframe = visual_odometry::Keyframe(tracker_->getPatches(), sample.timestamp);
				visualOdometry.newKeyframeCandidate(keyframe);
			}
			nTracksLast = tracker_->getPatches().size();
			if (dump)
				for (const auto& p : tracker_->getPatches())
					std::fprintf(dump, "%d %d %.17g %.17g\n", k, p.getTrackId(), p.toCorner().x, p.toCorner().y);
		}
		// ---- Evaluator::~Evaluator (evaluator.cpp:15-21)
		tracker_->preExit();
		nArchived = tracker_->getArchivedPatches().size();
		nOptimisations = tracker_->getOptimizedFinalCosts().size();
	}
	else
	{
		const auto counts = readVec<int64_t>(f);
		const auto ids = readVec<int32_t>(f);
		const auto xy = readVec<double>(f);
		std::fclose(f);
		size_t base = 0;
		for (int k = 0; k < nF; ++k)
		{
			tracker::Patches patches;
			for (int64_t i = 0; i < counts[k]; ++i)
			{
				tracker::Patch p(tracker::Corner(xy[2 * (base + i)], xy[2 * (base + i) + 1]), 12,
								 common::timestamp_t(frameTs[k]));
				p.setTrackId(ids[base + i]);
				patches.push_back(p);
			}
			base += static_cast<size_t>(counts[k]);
			imageNum++;
			if (imageNum > 2)
			{
				auto keyframe = visual_odometry::Keyframe(patches, common::timestamp_t(frameTs[k]));
				visualOdometry.newKeyframeCandidate(keyframe);
			}
			nTracksLast = patches.size();
		}
	}

	// all keyframes the front end holds, as newKeyframeCandidate lists them (visual_odometry.cpp:78-82)
	std::list<vo::Keyframe> pose = visualOdometry.getStoredFrames();
	for (const auto& kf : visualOdometry.getActiveFrames()) pose.emplace_back(kf.second);
	std::vector<common::Pose3d> ref;
	common::Pose3d zero;
	bool haveZero = false, allGt = true;
	for (const auto& kf : pose)
	{
		const auto g = visualOdometry.syncGtAndImage(kf.timestamp);
		if (!g.has_value())
		{
			allGt = false;
			break;
		}
		if (!haveZero)
		{
			zero = g.value();
			haveZero = true;
		}
		ref.push_back(zero.inverse() * g.value());
	}
	vo::ErrorMetricValue ate;
	if (allGt && pose.size() > 2) vo::align_cameras_sim3(ref, pose, &ate);
	if (dump) std::fclose(dump);
	FILE* out = std::fopen(argv[3], "w");
	if (!out) return 2;
	std::fprintf(out, "mode %s\nkeyframes %zu\nstored %zu\nactive %zu\nlandmarks %zu\ntracks_last_image %zu\n", mode.c_str(),
				 pose.size(), visualOdometry.getStoredFrames().size(), visualOdometry.getActiveFrames().size(),
				 visualOdometry.getMapLandmarks().landmarks.size(), nTracksLast);
	std::fprintf(out, "optimisations %zu\narchived %zu\n", nOptimisations, nArchived);
	std::fprintf(out, "ate_rmse %.17g\nate_mean %.17g\nate_max %.17g\nate_min %.17g\nate_count %.0f\n", ate.rmse, ate.mean,
				 ate.max, ate.min, ate.count);
	for (const auto& kf : pose)
		std::fprintf(out, "pose %lld %.17g %.17g %.17g\n", static_cast<long long>(kf.timestamp.count()),
					 kf.pose.translation().x(), kf.pose.translation().y(), kf.pose.translation().z());
	std::fclose(out);
	return 0;
}
