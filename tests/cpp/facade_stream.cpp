// Drives tracker::FeatureDetector (include/ebt_tracker.hpp) the way
// tools/evaluator/src/evaluator.cpp does — eventCallback -> updatePatches(event)
// one event at a time, imageCallback -> newImage — on a stream dumped by
// tests/test_gpu_facade.py, and prints the final tracks.
#include <cstdio>
#include <cstdlib>
#include <map>
#include <vector>

#include "ebt_tracker.hpp"

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
	if (argc < 2) return 2;
	FILE* f = std::fopen(argv[1], "rb");
	if (!f) return 2;
	const auto hdr = readVec<int64_t>(f);  // width, height, nFrames, maxPatches
	const int W = static_cast<int>(hdr[0]), H = static_cast<int>(hdr[1]), nF = static_cast<int>(hdr[2]);
	const auto frameTs = readVec<int64_t>(f);
	const auto frames = readVec<uint8_t>(f);
	const auto events = readVec<ebt_event>(f);
	const auto cornerCount = readVec<int64_t>(f);
	const auto corners = readVec<double>(f);
	const auto flows = readVec<double>(f);  // rows: frame, x, y, found, nx, ny
	std::fclose(f);

	tracker::DetectorParams params;
	params.imageWidth = W;
	params.imageHeight = H;
	params.maxPatches = static_cast<size_t>(hdr[3]);
	tracker::FeatureDetector det(params);
	int frame = 0;
	size_t cornerBase = 0;
	det.setCornerDetector([&](const uint8_t*, int, int) {
		tracker::Corners c;
		for (int64_t i = 0; i < cornerCount[frame]; ++i)
			c.push_back({corners[2 * (cornerBase + i)], corners[2 * (cornerBase + i) + 1]});
		return c;
	});
	det.setFlowTracker([&](const tracker::Corner& p, float& nx, float& ny) {
		for (size_t r = 0; r + 5 < flows.size() + 1; r += 6)
			if (static_cast<int>(flows[r]) == frame && flows[r + 1] == p.x && flows[r + 2] == p.y)
			{
				nx = static_cast<float>(flows[r + 4]);
				ny = static_cast<float>(flows[r + 5]);
				return flows[r + 3] != 0;
			}
		return false;
	});
	size_t e = 0;
	for (frame = 0; frame < nF; ++frame)
	{
		det.newImage(frames.data() + static_cast<size_t>(frame) * W * H, frameTs[frame]);
		cornerBase += static_cast<size_t>(cornerCount[frame]);
		const int64_t tEnd = frame + 1 < nF ? frameTs[frame + 1] : INT64_MAX;
		for (; e < events.size() && events[e].t_us < tEnd; ++e)
		{
			tracker::EventSample s;
			s.x = events[e].x;
			s.y = events[e].y;
			s.sign = events[e].polarity;
			s.timestamp = events[e].t_us;
			det.addEvent(s);
			det.updatePatches(s);  // evaluator.cpp:34-35
		}
	}
	for (const auto& p : det.getPatches())
		std::printf("T %d %d %.17g %.17g %.9g %zu %zu\n", p.getTrackId(), p.isLost() ? 1 : 0, p.toCorner().x,
					p.toCorner().y, static_cast<double>(p.getFlow()), p.getTrajectory().size(),
					p.getFinalCosts().size());
	std::printf("N %zu %zu %zu\n", det.getPatches().size(), det.getArchivedPatches().size(),
				det.getOptimizedFinalCosts().size());
	return 0;
}
