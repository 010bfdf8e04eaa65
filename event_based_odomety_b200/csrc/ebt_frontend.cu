// ebt_frontend.cu — sm_100a kernels of the per-frame re-arm:
// what FeatureDetector::newImage delegates to OpenCV in the reference.
//
//   detectFeatures  = cv::goodFeaturesToTrack(gray, W*H/625, 0.01, 10, mask, 3, useHarris, 0.04)
//                     (/root/reference/implementation/feature_tracker/src/feature_detector.cpp:568-583)
//   FlowEstimator::getFlow = cv::calcOpticalFlowPyrLK(prev, cur, {pt}, 21x21, maxLevel 3)
//                     (src/flow_estimator.cpp:84-103)
//
// OpenCV is a system dependency of the reference; the algorithms restated here
// are OpenCV 4.x's (imgproc/corner.cpp, imgproc/featureselect.cpp,
// video/lkpyramid.cpp), pinned against cv2 golden vectors
// (tests/golden/frontend_cv2.npz).  Byte/integer work (pyramid, Scharr,
// fixed-point bilinear patches) is bit-exact; float arithmetic is written
// without FMA contraction (explicit __f*_rn) in OpenCV's scalar order.
//
// All of it is once-per-frame, HBM/L2-bound work on <= 0.3 Mpx images.
#include <cuda_runtime.h>
#include <stdint.h>

#include "ebt_device_types.h"
#include "ebt_frontend.h"

namespace
{
__device__ __forceinline__ int reflect101(int i, int n)
{
	// BORDER_REFLECT_101: gfedcb|abcdefgh|gfedcba ; valid for -n < i < 2n-1
	if (i < 0) i = -i;
	if (i >= n) i = 2 * n - 2 - i;
	return i;
}

// ---------------------------------------------------------------------------
// Harris response (corner.cpp cornerEigenValsVecs + calcHarris), block 3, ksize 3
// ---------------------------------------------------------------------------
// pass 1: Sobel 3x3 (exact small integers) * scale, then the three products
__global__ void harris_products_kernel(const unsigned char* __restrict__ gray, int W, int H,
									   float* __restrict__ pa, float* __restrict__ pb,
									   float* __restrict__ pc)
{
	const int x = blockIdx.x * blockDim.x + threadIdx.x, y = blockIdx.y * blockDim.y + threadIdx.y;
	if (x >= W || y >= H) return;
	const int xm = reflect101(x - 1, W), xp = reflect101(x + 1, W);
	const int ym = reflect101(y - 1, H), yp = reflect101(y + 1, H);
	const float p00 = gray[ym * W + xm], p01 = gray[ym * W + x], p02 = gray[ym * W + xp];
	const float p10 = gray[y * W + xm], p12 = gray[y * W + xp];
	const float p20 = gray[yp * W + xm], p21 = gray[yp * W + x], p22 = gray[yp * W + xp];
	const float scale = (float)(1.0 / (4.0 * 3.0 * 255.0));
	const float dxi = (p12 - p10) * 2.0f + (p02 - p00) + (p22 - p20);  // exact: |.| <= 1020
	const float dyi = (p21 - p01) * 2.0f + (p20 - p00) + (p22 - p02);
	const float dx = __fmul_rn(dxi, scale), dy = __fmul_rn(dyi, scale);
	const int o = y * W + x;
	pa[o] = __fmul_rn(dx, dx);
	pb[o] = __fmul_rn(dx, dy);
	pc[o] = __fmul_rn(dy, dy);
}

__device__ __forceinline__ unsigned ordered_key(float f)
{
	const unsigned b = __float_as_uint(f);
	return (b & 0x80000000u) ? ~b : (b | 0x80000000u);
}
__device__ __forceinline__ float key_to_float(unsigned k)
{
	return __uint_as_float((k & 0x80000000u) ? (k & 0x7fffffffu) : ~k);
}

__device__ __forceinline__ float box3(const float* __restrict__ p, int x, int y, int W, int H)
{
	float s = 0.0f;
#pragma unroll
	for (int i = -1; i <= 1; ++i)
	{
		const int yy = reflect101(y + i, H);
#pragma unroll
		for (int j = -1; j <= 1; ++j) s = __fadd_rn(s, p[yy * W + reflect101(x + j, W)]);
	}
	return s;
}

// pass 2: un-normalised 3x3 box sums, response, global maximum
__global__ void harris_response_kernel(const float* __restrict__ pa, const float* __restrict__ pb,
									   const float* __restrict__ pc, int W, int H, float k,
									   float* __restrict__ resp, unsigned* __restrict__ max_key)
{
	const int x = blockIdx.x * blockDim.x + threadIdx.x, y = blockIdx.y * blockDim.y + threadIdx.y;
	unsigned key = 0;
	if (x < W && y < H)
	{
		const float a = box3(pa, x, y, W, H), b = box3(pb, x, y, W, H), c = box3(pc, x, y, W, H);
		const float tr = __fadd_rn(a, c);
		const float r = __fsub_rn(__fsub_rn(__fmul_rn(a, c), __fmul_rn(b, b)),
								  __fmul_rn(__fmul_rn(k, tr), tr));
		resp[y * W + x] = r;
		key = ordered_key(r);
	}
#pragma unroll
	for (int o = 16; o > 0; o >>= 1) key = max(key, __shfl_xor_sync(0xffffffffu, key, o));
	if (((threadIdx.y * blockDim.x + threadIdx.x) & 31) == 0 && key) atomicMax(max_key, key);
}

// pass 3: threshold at quality * max, 3x3 non-maximum suppression on the
// interior, mask = image minus an inset border; survivors are appended to a list
__global__ void harris_nms_kernel(const float* __restrict__ resp, int W, int H, const unsigned* max_key,
								  double quality, int inset, ebt_corner_candidate* __restrict__ out,
								  int cap, int* __restrict__ n_out)
{
	const int x = blockIdx.x * blockDim.x + threadIdx.x, y = blockIdx.y * blockDim.y + threadIdx.y;
	if (x < 1 || y < 1 || x >= W - 1 || y >= H - 1) return;
	if (x < inset || y < inset || x >= W - inset || y >= H - inset) return;
	// cv::threshold(eig, eig, maxVal * qualityLevel, 0, THRESH_TOZERO): double product, float compare
	const float thr = (float)((double)key_to_float(*max_key) * quality);
	const float v = resp[y * W + x];
	if (!(v > thr) || v == 0.0f) return;
	float m = v;
#pragma unroll
	for (int i = -1; i <= 1; ++i)
#pragma unroll
		for (int j = -1; j <= 1; ++j)
		{
			const float q = resp[(y + i) * W + (x + j)];
			m = fmaxf(m, q > thr ? q : 0.0f);
		}
	if (v != m) return;
	const int slot = atomicAdd(n_out, 1);
	if (slot < cap)
	{
		out[slot].value = v;
		out[slot].x = x;
		out[slot].y = y;
	}
}

// ---------------------------------------------------------------------------
// Pyramid and derivatives (lkpyramid.cpp / pyramids.cpp), bit-exact integer work
// ---------------------------------------------------------------------------
__global__ void pyr_down_kernel(const unsigned char* __restrict__ src, int W, int H,
								unsigned char* __restrict__ dst, int OW, int OH)
{
	const int x = blockIdx.x * blockDim.x + threadIdx.x, y = blockIdx.y * blockDim.y + threadIdx.y;
	if (x >= OW || y >= OH) return;
	const int kw[5] = {1, 4, 6, 4, 1};
	int acc = 0;
#pragma unroll
	for (int i = 0; i < 5; ++i)
	{
		const int yy = reflect101(2 * y - 2 + i, H);
		int row = 0;
#pragma unroll
		for (int j = 0; j < 5; ++j) row += kw[j] * (int)src[yy * W + reflect101(2 * x - 2 + j, W)];
		acc += kw[i] * row;
	}
	dst[y * OW + x] = (unsigned char)((acc + 128) >> 8);
}

// calcSharrDeriv: (dx, dy) int16, reflect-101 at the image border
__global__ void scharr_kernel(const unsigned char* __restrict__ src, int W, int H, short2* __restrict__ dst)
{
	const int x = blockIdx.x * blockDim.x + threadIdx.x, y = blockIdx.y * blockDim.y + threadIdx.y;
	if (x >= W || y >= H) return;
	const int ym = reflect101(y - 1, H), yp = reflect101(y + 1, H);
	int t0[3], t1[3];
#pragma unroll
	for (int j = 0; j < 3; ++j)
	{
		const int xx = reflect101(x - 1 + j, W);
		const int a = src[ym * W + xx], b = src[y * W + xx], c = src[yp * W + xx];
		t0[j] = (a + c) * 3 + b * 10;
		t1[j] = c - a;
	}
	dst[y * W + x] = make_short2((short)(t0[2] - t0[0]), (short)((t1[2] + t1[0]) * 3 + t1[1] * 10));
}

// ---------------------------------------------------------------------------
// LKTrackerInvoker (lkpyramid.cpp), one CTA per point, all levels in one launch
// ---------------------------------------------------------------------------
#define LK_WIN EBT_LK_WINDOW
#define LK_AREA (LK_WIN * LK_WIN)
#define LK_THREADS 128
#define LK_W_BITS 14

__device__ __forceinline__ int descale(int x, int n) { return (x + (1 << (n - 1))) >> n; }

// block sum of a double (LK_THREADS threads); every thread gets the total
__device__ double lk_block_sum(double v, double* scratch)
{
#pragma unroll
	for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
	__syncthreads();
	if ((threadIdx.x & 31) == 0) scratch[threadIdx.x >> 5] = v;
	__syncthreads();
	double t = 0;
#pragma unroll
	for (int w = 0; w < LK_THREADS / 32; ++w) t += scratch[w];
	return t;
}

__device__ __forceinline__ void lk_weights(float a, float b, int (&iw)[4])
{
	const float s = (float)(1 << LK_W_BITS);
	iw[0] = __float2int_rn(__fmul_rn(__fmul_rn(__fsub_rn(1.f, a), __fsub_rn(1.f, b)), s));
	iw[1] = __float2int_rn(__fmul_rn(__fmul_rn(a, __fsub_rn(1.f, b)), s));
	iw[2] = __float2int_rn(__fmul_rn(__fmul_rn(__fsub_rn(1.f, a), b), s));
	iw[3] = (1 << LK_W_BITS) - iw[0] - iw[1] - iw[2];
}

__global__ void __launch_bounds__(LK_THREADS)
lk_track_kernel(ebt_lk_pyramids pyr, const float* __restrict__ pts, int n_pts,
				float* __restrict__ next_pts, unsigned char* __restrict__ status, int max_count,
				double eps2)
{
	__shared__ short Iw[LK_AREA], Ix[LK_AREA], Iy[LK_AREA];
	__shared__ double scratch[LK_THREADS / 32];
	const int p = blockIdx.x;
	if (p >= n_pts) return;
	const float half = (LK_WIN - 1) * 0.5f;
	const float px0 = pts[2 * p], py0 = pts[2 * p + 1];
	float nx = 0.f, ny = 0.f;  // nextPts[ptidx]
	bool ok = true;
	for (int level = pyr.levels - 1; level >= 0; --level)
	{
		const int W = pyr.w[level], H = pyr.h[level];
		const unsigned char* I = pyr.prev[level];
		const unsigned char* J = pyr.cur[level];
		const short2* D = pyr.deriv[level];
		const float sc = (float)(1. / (1 << level));
		float ppx = __fmul_rn(px0, sc), ppy = __fmul_rn(py0, sc);
		float qx, qy;
		if (level == pyr.levels - 1)
		{
			qx = ppx;
			qy = ppy;
		}
		else
		{
			qx = __fmul_rn(nx, 2.f);
			qy = __fmul_rn(ny, 2.f);
		}
		nx = qx;
		ny = qy;
		ppx = __fsub_rn(ppx, half);
		ppy = __fsub_rn(ppy, half);
		const int ix = (int)floorf(ppx), iy = (int)floorf(ppy);
		if (ix < -LK_WIN || ix >= W || iy < -LK_WIN || iy >= H)
		{
			if (level == 0) ok = false;
			continue;
		}
		int iw[4];
		lk_weights(__fsub_rn(ppx, (float)ix), __fsub_rn(ppy, (float)iy), iw);
		double s11 = 0, s12 = 0, s22 = 0;
		__syncthreads();  // previous level's patches are no longer read
		for (int k = threadIdx.x; k < LK_AREA; k += LK_THREADS)
		{
			const int wy = k / LK_WIN, wx = k - wy * LK_WIN;
			const int y0 = iy + wy, x0 = ix + wx;
			const int ya = reflect101(y0, H), yb = reflect101(y0 + 1, H);
			const int xa = reflect101(x0, W), xb = reflect101(x0 + 1, W);
			const int iv = descale(I[ya * W + xa] * iw[0] + I[ya * W + xb] * iw[1] + I[yb * W + xa] * iw[2] +
									   I[yb * W + xb] * iw[3],
								   LK_W_BITS - 5);
			// derivative image: constant 0 outside (copyMakeBorder BORDER_CONSTANT)
			const bool r0 = y0 >= 0 && y0 < H, r1 = y0 + 1 >= 0 && y0 + 1 < H;
			const bool c0 = x0 >= 0 && x0 < W, c1 = x0 + 1 >= 0 && x0 + 1 < W;
			const short2 z = make_short2(0, 0);
			const short2 d00 = (r0 && c0) ? D[y0 * W + x0] : z, d01 = (r0 && c1) ? D[y0 * W + x0 + 1] : z;
			const short2 d10 = (r1 && c0) ? D[(y0 + 1) * W + x0] : z, d11 = (r1 && c1) ? D[(y0 + 1) * W + x0 + 1] : z;
			const int gx = descale(d00.x * iw[0] + d01.x * iw[1] + d10.x * iw[2] + d11.x * iw[3], LK_W_BITS);
			const int gy = descale(d00.y * iw[0] + d01.y * iw[1] + d10.y * iw[2] + d11.y * iw[3], LK_W_BITS);
			const short siv = (short)iv, sgx = (short)gx, sgy = (short)gy;
			Iw[k] = siv;
			Ix[k] = sgx;
			Iy[k] = sgy;
			s11 += (double)(float)((int)sgx * (int)sgx);
			s12 += (double)(float)((int)sgx * (int)sgy);
			s22 += (double)(float)((int)sgy * (int)sgy);
		}
		const float FLT_SCALE = 1.f / (1 << 20);
		const float A11 = __fmul_rn((float)lk_block_sum(s11, scratch), FLT_SCALE);
		const float A12 = __fmul_rn((float)lk_block_sum(s12, scratch), FLT_SCALE);
		const float A22 = __fmul_rn((float)lk_block_sum(s22, scratch), FLT_SCALE);
		float Dd = __fsub_rn(__fmul_rn(A11, A22), __fmul_rn(A12, A12));
		const float dif = __fsub_rn(A11, A22);
		const float disc = __fadd_rn(__fmul_rn(dif, dif), __fmul_rn(__fmul_rn(4.f, A12), A12));
		const float min_eig = __fdiv_rn(__fsub_rn(__fadd_rn(A22, A11), __fsqrt_rn(disc)), (float)(2 * LK_AREA));
		if (min_eig < 1e-4f || Dd < 1.1920929e-07f)
		{
			if (level == 0) ok = false;
			continue;
		}
		Dd = __fdiv_rn(1.f, Dd);
		qx = __fsub_rn(qx, half);
		qy = __fsub_rn(qy, half);
		float pdx = 0.f, pdy = 0.f;
		for (int j = 0; j < max_count; ++j)
		{
			const int jx = (int)floorf(qx), jy = (int)floorf(qy);
			if (jx < -LK_WIN || jx >= W || jy < -LK_WIN || jy >= H)
			{
				if (level == 0) ok = false;
				break;
			}
			lk_weights(__fsub_rn(qx, (float)jx), __fsub_rn(qy, (float)jy), iw);
			double t1 = 0, t2 = 0;
			for (int k = threadIdx.x; k < LK_AREA; k += LK_THREADS)
			{
				const int wy = k / LK_WIN, wx = k - wy * LK_WIN;
				const int ya = reflect101(jy + wy, H), yb = reflect101(jy + wy + 1, H);
				const int xa = reflect101(jx + wx, W), xb = reflect101(jx + wx + 1, W);
				const int jv = descale(J[ya * W + xa] * iw[0] + J[ya * W + xb] * iw[1] + J[yb * W + xa] * iw[2] +
										   J[yb * W + xb] * iw[3],
									   LK_W_BITS - 5);
				const int diff = jv - (int)Iw[k];
				t1 += (double)(float)(diff * (int)Ix[k]);
				t2 += (double)(float)(diff * (int)Iy[k]);
			}
			const float b1 = __fmul_rn((float)lk_block_sum(t1, scratch), FLT_SCALE);
			const float b2 = __fmul_rn((float)lk_block_sum(t2, scratch), FLT_SCALE);
			const float dx = __fmul_rn(__fsub_rn(__fmul_rn(A12, b2), __fmul_rn(A22, b1)), Dd);
			const float dy = __fmul_rn(__fsub_rn(__fmul_rn(A12, b1), __fmul_rn(A11, b2)), Dd);
			qx = __fadd_rn(qx, dx);
			qy = __fadd_rn(qy, dy);
			nx = __fadd_rn(qx, half);
			ny = __fadd_rn(qy, half);
			if ((double)dx * (double)dx + (double)dy * (double)dy <= eps2) break;
			if (j > 0 && (double)fabsf(__fadd_rn(dx, pdx)) < 0.01 && (double)fabsf(__fadd_rn(dy, pdy)) < 0.01)
			{
				nx = __fsub_rn(nx, __fmul_rn(dx, 0.5f));
				ny = __fsub_rn(ny, __fmul_rn(dy, 0.5f));
				break;
			}
			pdx = dx;
			pdy = dy;
		}
	}
	if (threadIdx.x == 0)
	{
		next_pts[2 * p] = nx;
		next_pts[2 * p + 1] = ny;
		status[p] = ok ? 1 : 0;
	}
}
// ---------------------------------------------------------------------------
// Tracker -> VO hand-off: Keyframe landmarks (vo/src/keyframe.cpp:5-14) and
// CameraModel::unproject (common/include/common/camera_model.h:79-114), one
// thread per landmark, FP64, the reference's operation order without FMA
// contraction.
// ---------------------------------------------------------------------------
__device__ __forceinline__ double tangential(double p1, double p2, double x, double y, double r2)
{
	// Scalar(2) * p1 * x * y + p2 * (r2 + Scalar(2) * x * x)
	return __dadd_rn(__dmul_rn(__dmul_rn(__dmul_rn(2.0, p1), x), y),
					 __dmul_rn(p2, __dadd_rn(r2, __dmul_rn(__dmul_rn(2.0, x), x))));
}

__global__ void unproject_kernel(const DevPatch* __restrict__ patches, const double* __restrict__ pts, int n,
								 ebt_camera cam, int do_bearings, int* __restrict__ ids,
								 double* __restrict__ corners, double* __restrict__ bearings)
{
	const int i = blockIdx.x * blockDim.x + threadIdx.x;
	if (i >= n) return;
	double x, y;
	if (patches)
	{
		x = patches[i].rx + (EBT_PATCH_SIDE - 1) / 2.;  // Patch::toCorner
		y = patches[i].ry + (EBT_PATCH_SIDE - 1) / 2.;
		ids[i] = patches[i].track_id;
		corners[2 * i] = x;
		corners[2 * i + 1] = y;
	}
	else
	{
		x = pts[2 * i];
		y = pts[2 * i + 1];
	}
	if (!do_bearings) return;
	const double xd = __ddiv_rn(x - cam.cx, cam.fx), yd = __ddiv_rn(y - cam.cy, cam.fy);
	double xo = xd, yo = yd;
	for (int it = 0; it < 10; ++it)
	{
		const double r2 = __dadd_rn(__dmul_rn(xo, xo), __dmul_rn(yo, yo));
		// 1 + k1 r2 + k2 r2 r2
		const double radial = __dadd_rn(__dadd_rn(1.0, __dmul_rn(cam.k1, r2)), __dmul_rn(__dmul_rn(cam.k2, r2), r2));
		const double dx = tangential(cam.p1, cam.p2, xo, yo, r2);
		const double dy = tangential(cam.p2, cam.p1, yo, xo, r2);
		xo = __ddiv_rn(xd - dx, radial);
		yo = __ddiv_rn(yd - dy, radial);
	}
	const double norm = __dsqrt_rn(__dadd_rn(__dadd_rn(__dmul_rn(xo, xo), __dmul_rn(yo, yo)), 1.0));
	bearings[3 * i] = __ddiv_rn(xo, norm);
	bearings[3 * i + 1] = __ddiv_rn(yo, norm);
	bearings[3 * i + 2] = __ddiv_rn(1.0, norm);
}
}  // namespace

// ---------------------------------------------------------------------------
// Launchers
// ---------------------------------------------------------------------------
static inline dim3 grid2d(int W, int H, dim3 b) { return dim3((W + b.x - 1) / b.x, (H + b.y - 1) / b.y); }

cudaError_t ebt_launch_harris(const unsigned char* gray, int W, int H, float k, double quality, int inset,
							  float* planes3, float* resp, unsigned* max_key, ebt_corner_candidate* cand,
							  int cap, int* n_cand, cudaStream_t stream)
{
	const dim3 b(32, 8);
	const dim3 g = grid2d(W, H, b);
	const size_t npx = (size_t)W * H;
	cudaError_t e = cudaMemsetAsync(max_key, 0, sizeof(unsigned), stream);
	if (e != cudaSuccess) return e;
	e = cudaMemsetAsync(n_cand, 0, sizeof(int), stream);
	if (e != cudaSuccess) return e;
	harris_products_kernel<<<g, b, 0, stream>>>(gray, W, H, planes3, planes3 + npx, planes3 + 2 * npx);
	harris_response_kernel<<<g, b, 0, stream>>>(planes3, planes3 + npx, planes3 + 2 * npx, W, H, k, resp, max_key);
	harris_nms_kernel<<<g, b, 0, stream>>>(resp, W, H, max_key, quality, inset, cand, cap, n_cand);
	return cudaGetLastError();
}

cudaError_t ebt_launch_pyr_down(const unsigned char* src, int W, int H, unsigned char* dst, cudaStream_t stream)
{
	const int OW = (W + 1) / 2, OH = (H + 1) / 2;
	const dim3 b(32, 8);
	pyr_down_kernel<<<grid2d(OW, OH, b), b, 0, stream>>>(src, W, H, dst, OW, OH);
	return cudaGetLastError();
}

cudaError_t ebt_launch_scharr(const unsigned char* src, int W, int H, short2* dst, cudaStream_t stream)
{
	const dim3 b(32, 8);
	scharr_kernel<<<grid2d(W, H, b), b, 0, stream>>>(src, W, H, dst);
	return cudaGetLastError();
}

cudaError_t ebt_launch_lk(const ebt_lk_pyramids& pyr, const float* pts, int n, float* next_pts,
						  unsigned char* status, int max_count, double eps, cudaStream_t stream)
{
	if (n <= 0) return cudaSuccess;
	lk_track_kernel<<<n, LK_THREADS, 0, stream>>>(pyr, pts, n, next_pts, status, max_count, eps * eps);
	return cudaGetLastError();
}

cudaError_t ebt_launch_unproject(const DevPatch* patches, const double* pts, int n, const ebt_camera& cam,
								 int do_bearings, int* ids, double* corners, double* bearings,
								 cudaStream_t stream)
{
	if (n <= 0) return cudaSuccess;
	unproject_kernel<<<(n + 127) / 128, 128, 0, stream>>>(patches, pts, n, cam, do_bearings, ids, corners, bearings);
	return cudaGetLastError();
}
