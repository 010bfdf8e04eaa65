// ebt_frontend.h — launchers of the per-frame re-arm kernels in ebt_frontend.cu
// (Harris corners, image pyramid, Scharr derivatives, pyramidal LK).
#pragma once
#include <cuda_runtime.h>

#define EBT_LK_MAX_LEVELS 4  // maxLevel 3 (FlowEstimatorParams::numLevels) + level 0
#define EBT_LK_WINDOW 21     // FlowEstimatorParams::window

struct ebt_corner_candidate
{
	float value;  // Harris response
	int x, y;
};

// Device pointers of the two pyramids of one LK call; level 0 = full resolution.
struct ebt_lk_pyramids
{
	int levels;
	int w[EBT_LK_MAX_LEVELS], h[EBT_LK_MAX_LEVELS];
	const unsigned char* prev[EBT_LK_MAX_LEVELS];
	const unsigned char* cur[EBT_LK_MAX_LEVELS];
	const short2* deriv[EBT_LK_MAX_LEVELS];  // Scharr (dx, dy) of prev
};

// Harris response + thresholded 3x3 local maxima inside the inset mask.
// planes3: 3*W*H floats scratch, resp: W*H floats; cand/n_cand: output list.
cudaError_t ebt_launch_harris(const unsigned char* gray, int W, int H, float k, double quality, int inset,
							  float* planes3, float* resp, unsigned* max_key, ebt_corner_candidate* cand,
							  int cap, int* n_cand, cudaStream_t stream);
cudaError_t ebt_launch_pyr_down(const unsigned char* src, int W, int H, unsigned char* dst, cudaStream_t stream);
cudaError_t ebt_launch_scharr(const unsigned char* src, int W, int H, short2* dst, cudaStream_t stream);
cudaError_t ebt_launch_lk(const ebt_lk_pyramids& pyr, const float* pts, int n, float* next_pts,
						  unsigned char* status, int max_count, double eps, cudaStream_t stream);

// CameraModel::unproject (common/include/common/camera_model.h:79-114) for n points.
// If patches != nullptr the points are the corners of patches[0..n) (Patch::toCorner) and
// ids / corners receive the Keyframe landmarks; otherwise points come from pts.
struct DevPatch;
struct ebt_camera;
cudaError_t ebt_launch_unproject(const DevPatch* patches, const double* pts, int n, const ebt_camera& cam,
								 int do_bearings, int* ids, double* corners, double* bearings,
								 cudaStream_t stream);
