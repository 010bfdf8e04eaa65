// TEST INFRASTRUCTURE ONLY (oracle/): BASELINE.json configs[0] — "feature_tracker via replayer on a
// shapes_6dof-format stream".  The reference's own dataset reader and replayer
// (tools/dataset_reader/src/{davis240c_reader,mapped_file}.cpp, tools/replayer/src/replayer.cpp), compiled
// UNMODIFIED, read a recording in the DAVIS240C text layout (events.txt, images.txt + image files) and
// deliver it through their callbacks, registered exactly as tools/evaluator/src/main.cpp registers
// Evaluator::eventCallback / imageCallback — into the drop-in tracker::FeatureDetector
// (include/ebt_dropin, CUDA library).  The output (optimisation records and final tracks) is compared by
// tests/test_replayer_c1.py with the library's own path for the same recording: ebt_parse_events_text
// (device parser of the same text) + ebt_replay (Replayer order).
//
//   replay_driver <dataset dir> <width> <height> <max patches> <out.txt>
#include <dataset_reader/davis240c_reader.h>
#include <feature_tracker/feature_detector.h>
#include <replayer/replayer.h>

#include <spdlog/sinks/null_sink.h>

#include <cstdio>
#include <cstdlib>
#include <memory>
#include <string>

int main(int argc, char** argv)
{
	if (argc < 6) return 2;
	auto sink = std::make_shared<spdlog::sinks::null_sink_mt>();
	spdlog::register_logger(std::make_shared<spdlog::logger>("console", sink));
	spdlog::register_logger(std::make_shared<spdlog::logger>("stderr", sink));

	tracker::DetectorParams params;
	params.imageSize = {std::atoi(argv[2]), std::atoi(argv[3])};
	params.maxPatches = static_cast<size_t>(std::atoi(argv[4]));
	tracker::FeatureDetector tracker_(params);

	auto reader = std::make_shared<tools::Davis240cReader>(argv[1]);
	tools::Replayer replayer(reader);
	size_t nEvents = 0, nImages = 0;
	// Evaluator::eventCallback (evaluator.cpp:32-35) and imageCallback (:69), tracker part
	replayer.addEventCallback([&](const common::EventSample& sample) {
		tracker_.addEvent(sample);
		tracker_.updatePatches(sample);
		++nEvents;
	});
	replayer.addImageCallback([&](const common::ImageSample& sample) {
		tracker_.newImage(sample);
		++nImages;
	});
	while (!replayer.finished()) replayer.next();

	FILE* out = std::fopen(argv[5], "w");
	if (!out) return 2;
	std::fprintf(out, "events %zu\nimages %zu\n", nEvents, nImages);
	for (const auto& r : tracker_.getOptimizedFinalCosts())
		std::fprintf(out, "rec %d %lld %.17g\n", r.trackId, static_cast<long long>(r.timeStampMicrosecond), r.lossValue);
	for (const auto& p : tracker_.getPatches())
		std::fprintf(out, "track %d %d %.17g %.17g %.9g %zu\n", p.getTrackId(), p.isLost() ? 1 : 0, p.toCorner().x,
					 p.toCorner().y, static_cast<double>(p.getFlow()), p.getTrajectory().size());
	for (const auto& p : tracker_.getArchivedPatches()) std::fprintf(out, "archived %d %zu\n", p.getTrackId(), p.getTrajectory().size());
	std::fclose(out);
	return 0;
}
