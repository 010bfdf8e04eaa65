// ebt_kernels.cu — hand-written sm_100a kernels for the event-based feature
// tracker hot path.  One CTA of EBT_THREADS threads owns
// one feature: its event window lives in shared memory, the 625 patch pixels
// are strided over the threads, per-evaluation sums are reduced with warp
// shuffles, and every thread carries an identical copy of the small scalar
// Levenberg-Marquardt state so that no broadcast is needed between steps.
//
// Arithmetic is FP64 throughout (B200 keeps a 1:2 FP64:FP32 vector rate), no
// tensor cores: the per-feature systems are 4x4.
//
// Reference behaviour mirrored (cited per function):
//   ft/     = /root/reference/implementation/feature_tracker/
//   ceres/  = /root/reference/thirdparty/ceres-solver/
//   sophus/ = /root/reference/thirdparty/Sophus/sophus/
#include <cuda_runtime.h>
#include <float.h>
#include <math.h>
#include <stdlib.h>

#include "ebt_device_types.h"
#include "ebt_kernels.h"

#define NPIX EBT_PATCH_PIXELS
#define SIDE EBT_PATCH_SIDE
#define NT EBT_THREADS
#define NW (EBT_THREADS / 32)
// EBT_DEDICATED_LEADER: warp 0 runs the scalar trust-region chain only and the
// pixel passes of the solve are shared by warps 1.. (PIX_NT threads), so the
// leader's LM state never has to survive a pixel pass in registers (or be
// spilled around it) and the kernel's register need is max(LM, pixel pass)
// instead of their sum.
#ifndef EBT_DEDICATED_LEADER
#define EBT_DEDICATED_LEADER 0
#endif
#define PIX_T0 (EBT_DEDICATED_LEADER ? 32 : 0)  // first thread that evaluates pixels in a solve
#define PIX_NT (NT - PIX_T0)
#define PIX_NW (PIX_NT / 32)
#define RMASK (EBT_RING - 1)
#ifdef EBT_INLINE_ALL
#define EBT_NOINLINE
#else
#define EBT_NOINLINE __noinline__
#endif
#ifndef EBT_PIX_UNROLL
#define EBT_PIX_UNROLL 1  // the hot loop must fit the 32 KB L1.5 instruction cache
#endif

// Phase clocks of the leader warp (build with -DEBT_PHASE_CLOCKS; read through
// ebt_launch_phase_clocks): 0 request+handshake, 1 pixel pass, 2 wait for the
// workers' partials, 3 totals+finalisation, 4 rest of the LM iteration,
// 5 evaluations counted.
__device__ unsigned long long g_phase_clocks[8];
#ifdef EBT_PHASE_CLOCKS
#define PHASE_T(v) const long long v = clock64()
#define PHASE_ADD(k, a, b) \
	if (threadIdx.x == 0) atomicAdd(&g_phase_clocks[k], (unsigned long long)((b) - (a)))
#else
#define PHASE_T(v)
#define PHASE_ADD(k, a, b)
#endif

// Walk clocks (build with -DEBT_WALK_CLOCKS, diagnostic builds only; they reuse the
// g_phase_clocks slots): summed over all features by the feature's first lane,
// 0 cycles entry->exit, 1 inside Optimizer::optimize, 2 inside lm_solve, 3 mask-row
// rebuilds, 4 rebuild count, 5 optimisations, 6 window-size update, 7 mask segments walked.
#ifdef EBT_WALK_CLOCKS
#define WALK_T(v) const long long v = clock64()
#define WALK_ADD(cond, k, amount) \
	if (cond) atomicAdd(&g_phase_clocks[k], (unsigned long long)(amount))
#else
#define WALK_T(v)
#define WALK_ADD(cond, k, amount)
#endif

namespace
{
constexpr int kPixUnroll = EBT_PIX_UNROLL;
// --------------------------------------------------------------------------
// Shared memory of one feature CTA
// --------------------------------------------------------------------------
// What one pixel evaluation needs besides the evaluation point
// Side of the square of template-grid texels staged in shared memory per solve:
// the warped 25x25 patch (+ 3 texels for the 4x4 taps) with ~3 texels of slack on
// every side for the motion of the iterates.
#define TS 36

struct EvalCtx
{
	const double2* grid;  // (gx, gy) template grid of the patch's frame
	int W, H;
	int r0, c0;           // grid texel of tile[0] (may lie outside the image: the tile is clamp-filled)
	int staged;           // tile holds grid[clamp(r0 + i)][clamp(c0 + j)]
	double tlx, tly;      // patch top-left
	double smm;           // sum m^2
	double huber_a;
};

struct __align__(16) FeatShared
{
	static constexpr int kPixWarps = (EBT_THREADS / 32) - (EBT_DEDICATED_LEADER ? 1 : 0);
	ebt_event ring[EBT_RING];  // Patch::events_
	double m[NPIX];            // normalised integrated patch (patch.cpp:156-159)
	int nabla[NPIX];           // Patch::integratedNabla_, integer valued
	double red[NW][20];        // cross-warp reduction scratch
	unsigned bal[2][NW];       // per-warp containment ballots (double buffered)
	double req[8];             // leader -> workers: evaluation point x[5], cos/sin(phi)
	double res[8];             // leader -> all: solved x[5], final cost
	int req_kind;              // 0 stop, 1 cost-only evaluation, 2 Jacobian evaluation
	int res_i[5];              // n_successful, n_unsuccessful, termination, n_jac, n_cost
	EvalCtx ctx;               // written by thread 0 before the solve
	double tot[20];            // CTA totals of one evaluation (leader warp)
	double2 tile[TS * TS];     // staged template-grid texels of the current solve (stage_tile)
	long long t_last;          // EBT_PHASE_CLOCKS only
};


__device__ __forceinline__ double warp_sum(double v)
{
#pragma unroll
	for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
	return v;
}

// Sum K per-thread partials over the CTA; every thread receives all K totals
// (fixed summation tree => bitwise identical in all threads).
//
// Warp stage: a transposing butterfly.  At each of the 5 shuffle steps a lane
// keeps one half of the values it still holds and hands the other half to its
// partner, so K values cost K/2 + K/4 + ... ~ K shuffles instead of 5 K; after
// the last step lane l holds the warp total of value slot_of_lane(l).
// Block stage: per-warp totals through shared memory, summed in warp order.
template <int K>
__device__ __forceinline__ int butterfly_slot(int lane)
{
	// mirrors Butterfly<K>::run: both halves are carried as H = ceil(n/2)
	// entries, the upper one zero padded when n is odd
	int lo = 0, n = K, real = K;
#pragma unroll
	for (int o = 16; o > 0; o >>= 1)
	{
		const int half = (n + 1) >> 1;
		if (lane & o)
		{
			lo += half;
			real -= half;
		}
		else
			real = min(real, half);
		n = half;
	}
	return real >= 1 ? lo : -1;
}

template <int N, int O>
struct Butterfly
{
	template <int K>
	static __device__ __forceinline__ void run(double (&v)[K], int lane)
	{
		constexpr int H = (N + 1) / 2;
		const bool up = (lane & O) != 0;
#pragma unroll
		for (int j = 0; j < H; ++j)
		{
			const double hi = (H + j < N) ? v[H + j] : 0.0;
			const double keep = up ? hi : v[j];
			const double give = up ? v[j] : hi;
			v[j] = keep + __shfl_xor_sync(0xffffffffu, give, O);
		}
		Butterfly<H, O / 2>::run(v, lane);
	}
};
template <int N>
struct Butterfly<N, 0>
{
	template <int K>
	static __device__ __forceinline__ void run(double (&)[K], int)
	{
	}
};

template <int K>
__device__ __forceinline__ void block_sum(FeatShared& sh, double (&acc)[K])
{
	const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
	Butterfly<K, 16>::run(acc, lane);
	const int slot = butterfly_slot<K>(lane);
	if (slot >= 0) sh.red[warp][slot] = acc[0];
	__syncthreads();
#pragma unroll
	for (int k = 0; k < K; ++k)
	{
		double v = sh.red[0][k];
#pragma unroll
		for (int w = 1; w < NW; ++w) v += sh.red[w][k];
		acc[k] = v;
	}
	__syncthreads();
}

// Catmull-Rom weights of ceres::CubicHermiteSpline (ceres/include/ceres/
// cubic_interpolation.h:64-90) in basis form: f = sum_k w[k] p[k],
// df/dx = sum_k d[k] p[k].
// The common factor 0.5 is left out here and applied once to the finished sums
// (0.25 for a row weight times a column weight): scaling by a power of two is
// exact, so the results are bit-identical and 10 FP64 multiplies per pixel go.
template <bool DERIV>
__device__ __forceinline__ void cr_weights(double t, double (&w)[4], double (&d)[4])
{
	const double t2 = t * t, t3 = t2 * t;
	w[0] = -t3 + 2.0 * t2 - t;
	w[1] = 3.0 * t3 - 5.0 * t2 + 2.0;
	w[2] = -3.0 * t3 + 4.0 * t2 + t;
	w[3] = t3 - t2;
	if (DERIV)
	{
		d[0] = -3.0 * t2 + 4.0 * t - 1.0;
		d[1] = 9.0 * t2 - 10.0 * t;
		d[2] = -9.0 * t2 + 8.0 * t + 1.0;
		d[3] = 3.0 * t2 - 2.0 * t;
	}
}

// Sum layout of one evaluation.
//  [0] sum u^2   [1] sum u*m   [2..5] A_q = sum u du_q   [6..9] B_q = sum du_q m
//  [10..19] G_qp = sum du_q du_p  (q <= p, row major upper triangle)
//
// Per pixel: OptimizerCostFunctor::warp (ft/include/feature_tracker/
// optimizer_cost.h:47-90): W = R p + t, out-of-image -> u = 0, else bicubic
// sample (ceres BiCubicInterpolator, cubic_interpolation.h:264-325, clamp to
// edge :403-421) of the interleaved (gx, gy) grid at (row = W.y, col = W.x),
// u = gx cos(phi) + gy sin(phi).  Jacobian columns: SE2 local (upsilon_x,
// upsilon_y, theta) through Dx_this_mul_exp_x_at_0 (sophus/se2.hpp:123-142)
// and phi.
// 4x4 neighbourhood of (row, col): INTERIOR = all 16 taps inside the image, so
// the clamp-to-edge index arithmetic (cubic_interpolation.h:403-421) is skipped
// and the taps are addressed from one base pointer.
template <bool JAC, bool INTERIOR>
__device__ __forceinline__ void sample_taps(const double2* __restrict__ grid, int W, int H, int row,
											int col, const double (&wr)[4], const double (&dr)[4],
											const double (&wc)[4], const double (&dc)[4], double& g0,
											double& g1, double& gr0, double& gr1, double& gc0,
											double& gc1)
{
	int ci[4];
	const double2* base = grid + ((row - 1) * W + (col - 1));  // 32-bit index: W*H < 2^31
	if (!INTERIOR)
	{
#pragma unroll
		for (int j = 0; j < 4; ++j) ci[j] = min(max(col - 1 + j, 0), W - 1);
	}
	g0 = g1 = gr0 = gr1 = gc0 = gc1 = 0;
#pragma unroll
	for (int r = 0; r < 4; ++r)
	{
		const double2* rowp =
			INTERIOR ? base + r * W : grid + min(max(row - 1 + r, 0), H - 1) * W;
		double f0 = 0, f1 = 0, d0 = 0, d1 = 0;
#pragma unroll
		for (int j = 0; j < 4; ++j)
		{
			const double2 p = __ldg(rowp + (INTERIOR ? j : ci[j]));
			f0 += wc[j] * p.x;
			f1 += wc[j] * p.y;
			if (JAC)
			{
				d0 += dc[j] * p.x;
				d1 += dc[j] * p.y;
			}
		}
		g0 += wr[r] * f0;
		g1 += wr[r] * f1;
		if (JAC)
		{
			gr0 += dr[r] * f0;
			gr1 += dr[r] * f1;
			gc0 += wr[r] * d0;
			gc1 += wr[r] * d1;
		}
	}
	// the two 0.5 factors cr_weights leaves out are applied by the caller (eval_pixels scales the flow
	// direction instead: a factor 0.25 is exact wherever it is applied)
}

// The same sums from the staged tile: `base` = texel (row - 1, col - 1) in shared
// memory, row stride TS.  The tile is filled with clamped indices, so it serves
// interior and border pixels alike; values and operation order are those of the
// global-memory form => bit-identical.
template <bool JAC>
__device__ __forceinline__ void sample_taps_tile(const double2* base, const double (&wr)[4],
												 const double (&dr)[4], const double (&wc)[4],
												 const double (&dc)[4], double& g0, double& g1, double& gr0,
												 double& gr1, double& gc0, double& gc1)
{
	g0 = g1 = gr0 = gr1 = gc0 = gc1 = 0;
#pragma unroll
	for (int r = 0; r < 4; ++r)
	{
		double f0 = 0, f1 = 0, d0 = 0, d1 = 0;
#pragma unroll
		for (int j = 0; j < 4; ++j)
		{
			const double2 p = base[r * TS + j];
			f0 += wc[j] * p.x;
			f1 += wc[j] * p.y;
			if (JAC)
			{
				d0 += dc[j] * p.x;
				d1 += dc[j] * p.y;
			}
		}
		g0 += wr[r] * f0;
		g1 += wr[r] * f1;
		if (JAC)
		{
			gr0 += dr[r] * f0;
			gr1 += dr[r] * f1;
			gc0 += wr[r] * d0;
			gc1 += wr[r] * d1;
		}
	}
}

#ifdef EBT_FP32_SAMPLING
// MEASUREMENT-ONLY variant (never the shipped build): the bicubic
// sampling, u and du of a pixel and the per-thread partial sums in FP32 (tile staged as float2),
// the reduction from the per-thread partials on, the in-image decisions and everything on the
// leader in FP64.  profiles/r2_variants.md reports its flip rate against the CPU reference and its speed.
template <bool DERIV>
__device__ __forceinline__ void cr_weights_f(float t, float (&w)[4], float (&d)[4])
{
	const float t2 = t * t, t3 = t2 * t;
	w[0] = -t3 + 2.0f * t2 - t;
	w[1] = 3.0f * t3 - 5.0f * t2 + 2.0f;
	w[2] = -3.0f * t3 + 4.0f * t2 + t;
	w[3] = t3 - t2;
	if (DERIV)
	{
		d[0] = -3.0f * t2 + 4.0f * t - 1.0f;
		d[1] = 9.0f * t2 - 10.0f * t;
		d[2] = -9.0f * t2 + 8.0f * t + 1.0f;
		d[3] = 3.0f * t2 - 2.0f * t;
	}
}
// stores a template-grid texel into the staged tile (float2 in this variant)
__device__ __forceinline__ void stage_texel(double2* tile, int idx, double2 g)
{
	reinterpret_cast<float2*>(tile)[idx] = make_float2((float)g.x, (float)g.y);
}
template <bool JAC, int STRIDE>
__device__ __forceinline__ void eval_pixels_f32(const double2* __restrict__ grid, int W, int H, const double (&x)[5],
												 double vx, double vy, double tlx, double tly, const double* m, int tid,
												 double (&acc)[JAC ? 20 : 2], const double2* tile, int r0, int c0,
												 bool staged)
{
	const double cs = x[0], sn = x[1], tx = x[2], ty = x[3];
	const float csf = (float)cs, snf = (float)sn, vxf = (float)vx, vyf = (float)vy;
	const float2* tf = reinterpret_cast<const float2*>(tile);
	float a[JAC ? 20 : 2];
#pragma unroll
	for (int k = 0; k < (JAC ? 20 : 2); ++k) a[k] = 0.0f;
#pragma unroll kPixUnroll
	for (int i = tid; i < NPIX; i += STRIDE)
	{
		const int yy = i / SIDE, xx = i - yy * SIDE;
		const double px = xx + tlx, py = yy + tly;
		const double wx = cs * px + -sn * py + tx;
		const double wy = sn * px + cs * py + ty;
		if (wx >= W || wy >= H || wx < 0.0 || wy < 0.0) continue;  // decision: FP64
		const double fr = floor(wy), fc = floor(wx);
		const int row = (int)fr, col = (int)fc;
		float wr[4], dr[4], wc[4], dc[4];
		cr_weights_f<JAC>((float)(wy - fr), wr, dr);
		cr_weights_f<JAC>((float)(wx - fc), wc, dc);
		float g0 = 0, g1 = 0, gr0 = 0, gr1 = 0, gc0 = 0, gc1 = 0;
		const int tr = row - 1 - r0, tc = col - 1 - c0;
		const bool in_tile = staged && tr >= 0 && tr + 3 < TS && tc >= 0 && tc + 3 < TS;
#pragma unroll
		for (int r = 0; r < 4; ++r)
		{
			float f0 = 0, f1 = 0, d0 = 0, d1 = 0;
#pragma unroll
			for (int j = 0; j < 4; ++j)
			{
				float2 p;
				if (in_tile)
					p = tf[(tr + r) * TS + tc + j];
				else
				{
					const double2 q = __ldg(grid + min(max(row - 1 + r, 0), H - 1) * W + min(max(col - 1 + j, 0), W - 1));
					p = make_float2((float)q.x, (float)q.y);
				}
				f0 += wc[j] * p.x;
				f1 += wc[j] * p.y;
				if (JAC)
				{
					d0 += dc[j] * p.x;
					d1 += dc[j] * p.y;
				}
			}
			g0 += wr[r] * f0;
			g1 += wr[r] * f1;
			if (JAC)
			{
				gr0 += dr[r] * f0;
				gr1 += dr[r] * f1;
				gc0 += wr[r] * d0;
				gc1 += wr[r] * d1;
			}
		}
		g0 *= 0.25f;
		g1 *= 0.25f;
		const float u = g0 * vxf + g1 * vyf;
		const float mi = (float)m[i];
		a[0] += u * u;
		a[1] += u * mi;
		if (JAC)
		{
			gr0 *= 0.25f; gr1 *= 0.25f; gc0 *= 0.25f; gc1 *= 0.25f;
			const float pxf = (float)px, pyf = (float)py;
			const float ux = gc0 * vxf + gc1 * vyf, uy = gr0 * vxf + gr1 * vyf;
			float du[4];
			du[0] = ux * csf + uy * snf;
			du[1] = -ux * snf + uy * csf;
			du[2] = ux * (-csf * pyf - snf * pxf) + uy * (-snf * pyf + csf * pxf);
			du[3] = -g0 * vyf + g1 * vxf;
#pragma unroll
			for (int q = 0; q < 4; ++q)
			{
				a[2 + q] += u * du[q];
				a[6 + q] += du[q] * mi;
			}
			int k = 10;
#pragma unroll
			for (int q = 0; q < 4; ++q)
#pragma unroll
				for (int pp = q; pp < 4; ++pp) a[k++] += du[q] * du[pp];
		}
	}
#pragma unroll
	for (int k = 0; k < (JAC ? 20 : 2); ++k) acc[k] = (double)a[k];
}
#else
__device__ __forceinline__ void stage_texel(double2* tile, int idx, double2 g) { tile[idx] = g; }
#endif

// The 625 pixels are strided over STRIDE threads; `tid` is this thread's index among them.
template <bool JAC, int STRIDE>
__device__ __forceinline__ void eval_pixels(const double2* __restrict__ grid, int W, int H,
											 const double (&x)[5], double vx, double vy,
											 double tlx, double tly, const double* m, int tid,
											 double (&acc)[JAC ? 20 : 2], const double2* tile, int r0, int c0,
											 bool staged)
{
#ifdef EBT_FP32_SAMPLING
	eval_pixels_f32<JAC, STRIDE>(grid, W, H, x, vx, vy, tlx, tly, m, tid, acc, tile, r0, c0, staged);
	return;
#endif
	const double cs = x[0], sn = x[1], tx = x[2], ty = x[3];
	// The samplers return 4x the interpolated values (the two 0.5 factors of the Catmull-Rom weights
	// are left out); everything below is linear in (vx, vy), so the factor goes onto the flow direction
	// once per pass instead of onto six sums per pixel.  A power of two is exact wherever it is
	// applied: the results are bit-identical.
	vx *= 0.25;
	vy *= 0.25;
#pragma unroll
	for (int k = 0; k < (JAC ? 20 : 2); ++k) acc[k] = 0.0;
#pragma unroll kPixUnroll
	for (int i = tid; i < NPIX; i += STRIDE)
	{
		const int yy = i / SIDE, xx = i - yy * SIDE;
		const double px = xx + tlx, py = yy + tly;
		const double wx = cs * px + -sn * py + tx;
		const double wy = sn * px + cs * py + ty;
		if (wx >= W || wy >= H || wx < 0.0 || wy < 0.0) continue;
		const double fr = floor(wy), fc = floor(wx);
		const int row = (int)fr, col = (int)fc;
		double wr[4], dr[4], wc[4], dc[4];
		cr_weights<JAC>(wy - fr, wr, dr);
		cr_weights<JAC>(wx - fc, wc, dc);
		double g0, g1, gr0, gr1, gc0, gc1;
		const int tr = row - 1 - r0, tc = col - 1 - c0;
		if (staged && tr >= 0 && tr + 3 < TS && tc >= 0 && tc + 3 < TS)
			sample_taps_tile<JAC>(tile + tr * TS + tc, wr, dr, wc, dc, g0, g1, gr0, gr1, gc0, gc1);
		else if (row >= 1 && row + 2 < H && col >= 1 && col + 2 < W)
			sample_taps<JAC, true>(grid, W, H, row, col, wr, dr, wc, dc, g0, g1, gr0, gr1, gc0, gc1);
		else
			sample_taps<JAC, false>(grid, W, H, row, col, wr, dr, wc, dc, g0, g1, gr0, gr1, gc0, gc1);
		const double u = g0 * vx + g1 * vy;
		const double mi = m[i];
		acc[0] += u * u;
		acc[1] += u * mi;
		if (JAC)
		{
			const double ux = gc0 * vx + gc1 * vy;  // du/dW.x
			const double uy = gr0 * vx + gr1 * vy;  // du/dW.y
			double du[4];
			du[0] = ux * cs + uy * sn;
			du[1] = -ux * sn + uy * cs;
			du[2] = ux * (-cs * py - sn * px) + uy * (-sn * py + cs * px);
			du[3] = -g0 * vy + g1 * vx;
#pragma unroll
			for (int q = 0; q < 4; ++q)
			{
				acc[2 + q] += u * du[q];
				acc[6 + q] += du[q] * mi;
			}
			int k = 10;
#pragma unroll
			for (int q = 0; q < 4; ++q)
#pragma unroll
				for (int p = q; p < 4; ++p) acc[k++] += du[q] * du[p];
		}
	}
}

// HuberLoss::Evaluate (ceres/internal/ceres/loss_function.cc:48-62) on the
// whole 625-block (ft/src/optimizer.cpp:97-99): returns rho(s), sets rho'.
__device__ __forceinline__ double huber(double s, double a, double& rho1)
{
	const double b = a * a;
	if (s > b)
	{
		const double r = sqrt(s);
		rho1 = fmax(DBL_MIN, a / r);
		return 2.0 * a * r - b;
	}
	rho1 = 1.0;
	return s;
}

// --------------------------------------------------------------------------
// Leader / worker split.  The scalar trust-region logic is one long dependent
// chain; it runs on warp 0 only (the "leader").  Warps 1..3 ("workers") sit in
// lm_worker() and join every pixel evaluation the leader posts:
//   leader : post request -> bar.sync 1 -> pixel_service -> bar.sync 2 -> totals
//   worker :                 bar.sync 1 -> pixel_service -> bar.arrive 2
// so a worker blocks once per evaluation.  pixel_service is the only copy of
// the two pixel loops in the kernel (one call site per role, not inlined).
// --------------------------------------------------------------------------
__device__ __forceinline__ void handshake()
{
	asm volatile("bar.sync 1, %0;" ::"n"(EBT_THREADS) : "memory");
}
__device__ __forceinline__ void partials_arrive()
{
	asm volatile("bar.arrive 2, %0;" ::"n"(EBT_THREADS) : "memory");
}
__device__ __forceinline__ void partials_wait()
{
	asm volatile("bar.sync 2, %0;" ::"n"(EBT_THREADS) : "memory");
}

// max / min as one compare + select for the leader's chain (fmax/fmin on doubles
// expand to ~9 instructions and ~25 cycles of dependent latency on sm_100a,
// measured with tools/micro/fp64_latency.cu).  `d` is the operand that is
// never NaN; like fmax/fmin the result is `d` when `v` is NaN.
__device__ __forceinline__ double sel_max(double d, double v) { return v > d ? v : d; }
__device__ __forceinline__ double sel_min(double d, double v) { return v < d ? v : d; }

// Branch-free reciprocal / reciprocal square root for the leader's chain:
// MUFU seed (>= 20 bits) + two Newton steps, ~1 ulp, no slow-path call.  The
// arguments here are normal, finite and non-zero by construction.
__device__ __forceinline__ double rcp_fast(double x)
{
	double r;
	asm("rcp.approx.ftz.f64 %0, %1;" : "=d"(r) : "d"(x));
	double e = fma(-x, r, 1.0);
	r = fma(r, e, r);
	e = fma(-x, r, 1.0);
	return fma(r, e, r);
}
__device__ __forceinline__ double rsqrt_fast(double x)
{
	double y;
	asm("rsqrt.approx.ftz.f64 %0, %1;" : "=d"(y) : "d"(x));
	double e = fma(-x * y, y, 1.0);
	y = fma(0.5 * y, e, y);
	e = fma(-x * y, y, 1.0);
	return fma(0.5 * y, e, y);
}
// sqrt for x >= 0 (0 -> 0): x * rsqrt(x) with one Heron correction
__device__ __forceinline__ double sqrt_fast(double x)
{
	const double y = rsqrt_fast(sel_max(DBL_MIN, x));
	const double s = x * y;
	return fma(fma(-s, s, x), 0.5 * y, s);
}

// Paths that are (almost) never taken on real data are kept out of line.  Inlined, each of them sits in
// the middle of the leader's trust-region loop with a library call inside; with the three below behind
// calls the c3 stream kernel runs 5 % faster (60.2 -> 57.0 ms per step, profiles/r2_variants.md 7).
#ifndef EBT_COLD
#define EBT_COLD __noinline__
#endif

// sin and cos for the leader's chain: Cody-Waite reduction by pi/2 (three FMA
// terms, exact products) and the fdlibm kernel polynomials on [-pi/4, pi/4];
// < 1.5 ulp, no slow-path call for |x| < 1e6 (angles here are a few radians).
// (the library routine for huge / non-finite arguments stays out of line: it is never taken
// on real data and would otherwise be copied into each of the five places this inlines)
__device__ EBT_COLD void sincos_huge(double x, double* sn, double* cs) { sincos(x, sn, cs); }
__device__ __forceinline__ void sincos_fast(double x, double* sn, double* cs)
{
	if (!(fabs(x) < 1e6))
	{
		sincos_huge(x, sn, cs);
		return;
	}
	const double q = rint(x * 0.6366197723675814);
	double r = fma(-q, 1.5707963267948966, x);
	r = fma(-q, 6.123233995736766e-17, r);
	r = fma(-q, -1.4973849048591698e-33, r);
	const int n = (int)q;
	const double z = r * r;
	double ps = 1.58969099521155010221e-10;
	ps = fma(ps, z, -2.50507602534068634195e-08);
	ps = fma(ps, z, 2.75573137070700676789e-06);
	ps = fma(ps, z, -1.98412698298579493134e-04);
	ps = fma(ps, z, 8.33333333332248946124e-03);
	ps = fma(ps, z, -1.66666666666666324348e-01);
	double pc = -1.13596475577881948265e-11;
	pc = fma(pc, z, 2.08757232129817482790e-09);
	pc = fma(pc, z, -2.75573143513906633035e-07);
	pc = fma(pc, z, 2.48015872894767294178e-05);
	pc = fma(pc, z, -1.38888888888741095749e-03);
	pc = fma(pc, z, 4.16666666666666019037e-02);
	const double s0 = fma(r * z, ps, r);
	const double c0 = fma(z * z, pc, fma(-0.5, z, 1.0));
	const double sv = (n & 1) ? c0 : s0;
	const double cv = (n & 1) ? s0 : c0;
	*sn = (n & 2) ? -sv : sv;
	*cs = ((n + 1) & 2) ? -cv : cv;
}

// fmod(phi, 2 pi) (optimizer.cpp:170): for |phi| < 2 pi — every solve but a handful — fmod returns
// its argument; the library loop for the rest stays out of line.
__device__ EBT_COLD double wrap_flow_cold(double phi) { return fmod(phi, 2 * 3.14159265358979323846); }
__device__ __forceinline__ double wrap_flow(double phi)
{
	return fabs(phi) < 2 * 3.14159265358979323846 ? phi : wrap_flow_cold(phi);
}

__device__ __forceinline__ void post_request(FeatShared& sh, int kind, const double (&x)[5], double vx,
											 double vy)
{
	if ((threadIdx.x & 31) == 0)
	{
#pragma unroll
		for (int k = 0; k < 5; ++k) sh.req[k] = x[k];
		sh.req[5] = vx;
		sh.req[6] = vy;
		sh.req_kind = kind;
	}
	handshake();
}

// Cooperative part of one evaluation: every warp evaluates its pixels of the
// posted request and leaves its warp totals in sh.red[warp][].
__device__ __noinline__ void pixel_service(FeatShared& sh)
{
	const int tid = (int)threadIdx.x - PIX_T0;
	const int lane = tid & 31, warp = tid >> 5;
	double x[5];
#pragma unroll
	for (int k = 0; k < 5; ++k) x[k] = sh.req[k];
	const double vx = sh.req[5], vy = sh.req[6];
	const EvalCtx& E = sh.ctx;
	if (sh.req_kind == 2)
	{
		double acc[20];
		eval_pixels<true, PIX_NT>(E.grid, E.W, E.H, x, vx, vy, E.tlx, E.tly, sh.m, tid, acc, sh.tile, E.r0, E.c0,
								  E.staged != 0);
		Butterfly<20, 16>::run(acc, lane);
		const int slot = butterfly_slot<20>(lane);
		if (slot >= 0) sh.red[warp][slot] = acc[0];
	}
	else
	{
		double acc[2];
		eval_pixels<false, PIX_NT>(E.grid, E.W, E.H, x, vx, vy, E.tlx, E.tly, sh.m, tid, acc, sh.tile, E.r0, E.c0,
								   E.staged != 0);
		Butterfly<2, 16>::run(acc, lane);
		const int slot = butterfly_slot<2>(lane);
		if (slot >= 0) sh.red[warp][slot] = acc[0];
	}
}

__device__ __forceinline__ void leader_pixels(FeatShared& sh) { pixel_service(sh); }

__device__ void lm_worker(FeatShared& sh)
{
	for (;;)
	{
		handshake();
		if (sh.req_kind == 0) break;
		pixel_service(sh);
		partials_arrive();
	}
}

// CTA total of reduced value k: pairwise over the per-warp partials
template <class SH>
__device__ __forceinline__ double red_total(const SH& sh, int k)
{
	constexpr int PW = SH::kPixWarps;
	static_assert((PW & (PW - 1)) == 0, "pairwise tree over the pixel warps");
	double v[PW];
#pragma unroll
	for (int w = 0; w < PW; ++w) v[w] = sh.red[w][k];
#pragma unroll
	for (int n = PW; n > 1; n >>= 1)
#pragma unroll
		for (int w = 0; w < n / 2; ++w) v[w] = v[2 * w] + v[2 * w + 1];
	return v[0];
}

// Leader side of one evaluation at `x`.  kind 2: cost, H = J^T J (upper 4x4) and
// g = J^T r of the Huber-corrected 625-block; kind 1: cost only.
//   r_i = u_i/sqrt(N) + m_i, N = 1e-5 + sum u^2   (optimizer_cost.h:29-45)
//   dr_i/dq = du_iq/sqrt(N) - u_i A_q / N^1.5
// Loss correction: HuberLoss::Evaluate (loss_function.cc:48-62) on the block's
// squared norm; with rho'' <= 0 residuals and Jacobian are scaled by
// sqrt(rho') (residual_block.cc:134-175, corrector.cc:81-110).
// Returns false when the evaluation is not finite (ResidualBlock::Evaluate's
// IsArrayValid check).
struct NoOverlap
{
	__device__ __forceinline__ void operator()() const {}
};
// `overlap` is scalar work of the leader that does not depend on this
// evaluation's result; it runs while the pixel warps are busy.
__device__ __forceinline__ void wait_partials(FeatShared&) { partials_wait(); }
template <class SH, class Overlap = NoOverlap>
__device__ __forceinline__ bool leader_eval(SH& sh, int kind, const double (&x)[5],
											double& cost, double (&Hm)[10], double (&g)[4],
											Overlap overlap = Overlap())
{
	PHASE_T(t0);
	double vy, vx;
	sincos_fast(x[4], &vy, &vx);
	post_request(sh, kind, x, vx, vy);
	PHASE_T(t1);
#if !EBT_DEDICATED_LEADER
	leader_pixels(sh);
#endif
	overlap();
	PHASE_T(t2);
	wait_partials(sh);
	PHASE_T(t3);
	PHASE_ADD(0, t0, t1);
	PHASE_ADD(1, t1, t2);
	PHASE_ADD(2, t2, t3);
#ifdef EBT_PHASE_CLOCKS
	if (threadIdx.x == 0)
	{
		atomicAdd(&g_phase_clocks[5], 1ull);
		if (sh.t_last) atomicAdd(&g_phase_clocks[4], (unsigned long long)(t0 - sh.t_last));
	}
#endif
	const EvalCtx& E = sh.ctx;
	{
		// every leader lane that owns a reduction slot sums that slot's per-warp
		// partials; the totals are then shared through sh.tot
		const int lane = threadIdx.x & 31;
		const int slot = kind == 2 ? butterfly_slot<20>(lane) : butterfly_slot<2>(lane);
		if (slot >= 0) sh.tot[slot] = red_total(sh, slot);
		__syncwarp();
	}
	const double suu = sh.tot[0];
	const double c = sh.tot[1];
	const double N = 1e-5 + suu;
	const double iN = rcp_fast(N), isq = rsqrt_fast(N);
	const double s = suu * iN + 2.0 * c * isq + E.smm;
	// Huber
	const double a = E.huber_a, b = a * a;
	double rho0 = s, rho1 = 1.0;
	if (s > b)
	{
		const double ir = rsqrt_fast(s);
		rho0 = 2.0 * a * (s * ir) - b;
		rho1 = sel_max(DBL_MIN, a * ir);
	}
	cost = 0.5 * rho0;
	bool ok = isfinite(s);
	if (kind == 2)
	{
		double A[4], B[4];
#pragma unroll
		for (int q = 0; q < 4; ++q)
		{
			A[q] = sh.tot[2 + q];
			B[q] = sh.tot[6 + q];
		}
		// g_q = rho' [ A_q/N (1 - suu/N - c/sqrt(N)) + B_q/sqrt(N) ]
		const double k1 = rho1 * iN * (1.0 - suu * iN - c * isq);
		const double kb = rho1 * isq;
		const double k2 = rho1 * iN * iN * (2.0 - suu * iN);
		const double kg = rho1 * iN;
#pragma unroll
		for (int q = 0; q < 4; ++q) g[q] = A[q] * k1 + B[q] * kb;
		int k = 0;
#pragma unroll
		for (int q = 0; q < 4; ++q)
#pragma unroll
			for (int p = q; p < 4; ++p, ++k)
			{
				const double G = sh.tot[10 + k];
				Hm[k] = G * kg - A[q] * A[p] * k2;
			}
		// any non-finite term poisons the sum (summed as a shallow tree: this is on
		// the leader's dependent chain)
		const double c0 = (g[0] + g[1]) + (g[2] + g[3]);
		const double c1 = (Hm[0] + Hm[1]) + (Hm[2] + Hm[3]);
		const double c2 = (Hm[4] + Hm[5]) + (Hm[6] + Hm[7]);
		const double c3 = Hm[8] + Hm[9];
		ok = ok && isfinite((c0 + c1) + (c2 + c3));
	}
#ifdef EBT_PHASE_CLOCKS
	{
		const long long t4 = clock64();
		PHASE_ADD(3, t3, t4);
		if (threadIdx.x == 0) sh.t_last = t4;
	}
#endif
	return ok;
}

// 1/|z| for z = (c, s) with |z|^2 = n2; unit up to rounding in every call here,
// where the first-order term is exact to double precision.
__device__ __forceinline__ double inv_len(double n2)
{
	const double e = n2 - 1.0;
	if (fabs(e) < 1e-7) return fma(e, fma(0.375, e, -0.5), 1.0);
	return rsqrt_fast(n2);
}

// LocalParameterizationSE2::Plus = T * SE2::exp(delta) (ft/include/
// feature_tracker/local_parameterization_se2.hpp:20-29; sophus/se2.hpp:560-582;
// SO2 product + renormalisation sophus/so2.hpp:214-239, normalising ctor
// :388-391); the flow angle block is plain addition.
__device__ __forceinline__ void plus5(const double (&x)[5], const double (&d)[4], double (&out)[5])
{
	const double theta = d[2];
	double es, ec;
	sincos_fast(theta, &es, &ec);
	{
		const double il = inv_len(ec * ec + es * es);
		ec *= il;
		es *= il;
	}
	double sbt, omc;
	if (fabs(theta) < 1e-10)
	{
		const double t2 = theta * theta;
		sbt = 1. - (1. / 6.) * t2;
		omc = 0.5 * theta - (1. / 24.) * theta * t2;
	}
	else
	{
		const double it = rcp_fast(theta);
		sbt = es * it;
		omc = (1. - ec) * it;
	}
	const double ex = sbt * d[0] - omc * d[1];
	const double ey = omc * d[0] + sbt * d[1];
	double rc = x[0] * ec - x[1] * es;
	double rs = x[0] * es + x[1] * ec;
	const double n2 = rc * rc + rs * rs;
	if (n2 != 1.0)
	{
		const double sc = 2.0 * rcp_fast(1.0 + n2);
		rc *= sc;
		rs *= sc;
	}
	{
		const double il = inv_len(rc * rc + rs * rs);
		rc *= il;
		rs *= il;
	}
	out[0] = rc;
	out[1] = rs;
	out[2] = x[2] + (x[0] * ex - x[1] * ey);
	out[3] = x[3] + (x[1] * ex + x[0] * ey);
	out[4] = x[4] + d[3];
}

__device__ __forceinline__ int hidx(int q, int p)
{
	// index of (q,p), q <= p, in the packed upper triangle of a 4x4
	return q * 4 - (q * (q - 1)) / 2 + (p - q);
}

// Solve (Hs + diag(d2)) y = gs by an LDL^T factorisation, FP64.  This is the
// normal-equation form of the least-squares problem DenseQRSolver solves by
// Householder QR (ceres/internal/ceres/dense_qr_solver.cc:120-165):
//   min || [J S; diag(sqrt(d2))] y - [r; 0] ||.
// Returns false when a pivot is not positive or y is not finite (Ceres:
// "Failed to compute a finite step", levenberg_marquardt_strategy.cc:117-121).
__device__ __forceinline__ bool solve4(const double (&Hs)[10], const double (&d2)[4],
									   const double (&gs)[4], double (&y)[4])
{
	double L[4][4], D[4], id[4];  // unit lower factor, pivots and their inverses
	bool pd = true;
#pragma unroll
	for (int j = 0; j < 4; ++j)
	{
		double w[4];
		double v = Hs[hidx(j, j)] + d2[j];
#pragma unroll
		for (int k = 0; k < j; ++k)
		{
			w[k] = L[j][k] * D[k];
			v -= L[j][k] * w[k];
		}
		D[j] = v;
		pd = pd && (v > 0.0);
		id[j] = rcp_fast(v);
#pragma unroll
		for (int i = j + 1; i < 4; ++i)
		{
			double t = Hs[hidx(j, i)];
#pragma unroll
			for (int k = 0; k < j; ++k) t -= L[i][k] * w[k];
			L[i][j] = t * id[j];
		}
	}
	double z[4];
#pragma unroll
	for (int i = 0; i < 4; ++i)
	{
		double v = gs[i];
#pragma unroll
		for (int k = 0; k < i; ++k) v -= L[i][k] * z[k];
		z[i] = v;
	}
#pragma unroll
	for (int i = 3; i >= 0; --i)
	{
		double v = z[i] * id[i];
#pragma unroll
		for (int k = i + 1; k < 4; ++k) v -= L[k][i] * y[k];
		y[i] = v;
	}
	return pd && isfinite((y[0] + y[1]) + (y[2] + y[3]));
}

struct LmSummary
{
	double final_cost;
	int n_successful, n_unsuccessful, termination;
	int n_jac, n_cost;
	int filtered;  // the cost-median filter dropped this solve (optimizer.cpp:151-156)
};

__device__ __forceinline__ double norm5(const double (&a)[5])
{
	double s = 0;
#pragma unroll
	for (int k = 0; k < 5; ++k) s += a[k] * a[k];
	return sqrt_fast(s);
}

// max |x - Plus(x, -g)| in full.  Out of line (EBT_COLD): the lower bound in scale_system decides
// the gradient-tolerance test on all but a handful of iterations, and inlined this is ~260
// instructions (a Plus with its sincos) in the middle of the trust-region loop, twice.
__device__ EBT_COLD double projected_gradient_norm(double x0, double x1, double x2, double x3, double x4, double g0,
												   double g1, double g2, double g3)
{
	const double x[5] = {x0, x1, x2, x3, x4};
	const double ng[4] = {-g0, -g1, -g2, -g3};
	double xp[5];
	plus5(x, ng, xp);
	double mx = 0;
#pragma unroll
	for (int k = 0; k < 5; ++k) mx = sel_max(mx, fabs(x[k] - xp[k]));
	return mx;
}

// Scaled system of EvaluateGradientAndJacobian (ceres/internal/ceres/
// trust_region_minimizer.cc:231-288): Jacobi column scaling fixed at iteration
// 0, Hs = S H S, gs = S g, and the projected gradient max norm
// max |x - Plus(x, -g)|.  The norm only feeds the gradient-tolerance test; its
// flow-angle component is |g[3]| up to one rounding of x[4], so the Plus is
// evaluated only when that lower bound does not already decide the test.
__device__ __forceinline__ void scale_system(const double (&x)[5], bool iter0, const double (&Hm)[10],
											 const double (&g)[4], double (&scale)[4],
											 double (&Hs)[10], double (&gs)[4], double tol,
											 double& grad_max_norm)
{
	if (iter0)
	{
#pragma unroll
		for (int q = 0; q < 4; ++q) scale[q] = rcp_fast(1.0 + sqrt_fast(sel_max(0.0, Hm[hidx(q, q)])));
	}
#pragma unroll
	for (int q = 0; q < 4; ++q)
	{
		gs[q] = g[q] * scale[q];
#pragma unroll
		for (int p = q; p < 4; ++p) Hs[hidx(q, p)] = Hm[hidx(q, p)] * scale[q] * scale[p];
	}
	const double lower = fabs(g[3]) - 8.0 * DBL_EPSILON * (fabs(x[4]) + fabs(g[3]));
	if (lower > tol)
	{
		grad_max_norm = fabs(g[3]);
		return;
	}
	grad_max_norm = projected_gradient_norm(x[0], x[1], x[2], x[3], x[4], g[0], g[1], g[2], g[3]);
}

// TrustRegionMinimizer::Minimize (trust_region_minimizer.cc:67-121 and helpers
// :300-344, 364-433, 438-470, 680-801), LevenbergMarquardtStrategy
// (levenberg_marquardt_strategy.cc:67-166), TrustRegionStepEvaluator
// (trust_region_step_evaluator.cc:40-114), final cost (solver_utils.h:48-56),
// with the reference's options (ft/src/optimizer.cpp:102-110).  Leader warp
// only.  One pixel evaluation per trust-region step: the candidate's cost and
// its Jacobian sums come from the same pass (Ceres evaluates the cost first and
// the Jacobian again after acceptance - same values, twice the pixel work);
// the sums are dropped when the step is rejected.  The step that reaches the
// iteration limit needs no Jacobian (the limit is tested before the gradient),
// so it is evaluated cost-only.  params is updated to the lowest-cost accepted
// iterate, or left alone on FAILURE (solver.cc:428-432).
template <class SH>
__device__ void lm_solve(SH& sh, const DevParams& P, double (&params)[5], LmSummary& S)
{
	double x[5], cand[5], best[5];
	double Hm[10], g[4];
	double Hs[10], gs[4], scale[4] = {1, 1, 1, 1}, diag[4] = {0, 0, 0, 0};
	double x_cost = DBL_MAX, minimum_cost = DBL_MAX, x_norm = -1;
	// decrease_factor_ of the trust-region strategy is 2, 4, 8, ...: its reciprocal is kept instead, so
	// that radius / decrease_factor is one exact multiplication (same bits: a power of two, no underflow
	// above min_radius) and no IEEE division sits in the loop
	double radius = P.initial_radius, inv_decrease = 0.5;
	bool reuse_diagonal = false;
	int consecutive_invalid = 0;
	double grad_max_norm = 0, final_cost = DBL_MAX;
	double se_min, se_cur, se_ref, se_cand, se_acc_ref = 0, se_acc_cand = 0;
	int se_nonmono = 0;
	int iteration = 0;
	bool failed = false;
#pragma unroll
	for (int k = 0; k < 5; ++k) x[k] = best[k] = params[k];
#ifdef EBT_PHASE_CLOCKS
	if (threadIdx.x == 0) sh.t_last = 0;
#endif
	S.n_successful = S.n_unsuccessful = S.termination = 0;
	S.n_jac = 1;
	S.n_cost = 0;
	if (!leader_eval(sh, 2, x, x_cost, Hm, g))
	{
		S.termination = 2;
		S.final_cost = -1.0;  // Summary::initial_cost default (solver.h)
		return;
	}
	scale_system(x, true, Hm, g, scale, Hs, gs, P.gradient_tolerance, grad_max_norm);
	final_cost = x_cost;
	se_min = se_cur = se_ref = se_cand = x_cost;  // TrustRegionStepEvaluator ctor
	bool step_successful = true;
	double iter_cost = x_cost, iter_gmn = grad_max_norm;
	for (;;)
	{
		// FinalizeIterationAndCheckIfMinimizerCanContinue
		if (step_successful)
		{
			if (iteration > 0) S.n_successful++;
			if (x_cost < minimum_cost)
			{
				minimum_cost = x_cost;
#pragma unroll
				for (int k = 0; k < 5; ++k) best[k] = x[k];
			}
		}
		else
			S.n_unsuccessful++;
		final_cost = sel_min(final_cost, iter_cost);
		if (iteration >= P.max_iterations) break;
		if (step_successful && iter_gmn <= P.gradient_tolerance) { S.termination = 1; break; }
		if (radius <= P.min_radius) { S.termination = 1; break; }
		++iteration;
		step_successful = false;
		iter_gmn = 0;
		iter_cost = 0;

		// LevenbergMarquardtStrategy::ComputeStep
		if (!reuse_diagonal)
		{
#pragma unroll
			for (int q = 0; q < 4; ++q)
				diag[q] = sel_min(P.max_lm_diagonal, sel_max(P.min_lm_diagonal, Hs[hidx(q, q)]));
		}
		double d2[4], step[4], delta[4];
		{
			const double ir = rcp_fast(radius);
#pragma unroll
			for (int q = 0; q < 4; ++q) d2[q] = diag[q] * ir;  // lm_diagonal_^2
		}
		bool valid = solve4(Hs, d2, gs, step);
		reuse_diagonal = true;
		double model_cost_change = 0;
		if (valid)
		{
			// -(J step)'(r + J step / 2) = -step'gs - step'Hs step / 2
			double sg = 0, shs = 0;
#pragma unroll
			for (int q = 0; q < 4; ++q)
			{
				step[q] = -step[q];
				sg += step[q] * gs[q];
			}
#pragma unroll
			for (int q = 0; q < 4; ++q)
			{
				double row = 0;
#pragma unroll
				for (int p = 0; p < 4; ++p) row += Hs[q <= p ? hidx(q, p) : hidx(p, q)] * step[p];
				shs += step[q] * row;
			}
			model_cost_change = -sg - 0.5 * shs;
			valid = model_cost_change > 0.0;
		}
		if (!valid)
		{
			// HandleInvalidStep
			if (++consecutive_invalid >= P.max_consecutive_invalid_steps)
			{
				S.termination = 2;
				failed = true;
				break;
			}
			radius *= inv_decrease;
			inv_decrease *= 0.5;
			iter_cost = x_cost;
			continue;
		}
		consecutive_invalid = 0;
#pragma unroll
		for (int q = 0; q < 4; ++q) delta[q] = step[q] * scale[q];
		plus5(x, delta, cand);
		const bool last = iteration >= P.max_iterations;
		double cand_cost;
		if (last)
			S.n_cost++;
		else
			S.n_jac++;
		// |x - cand| (parameter tolerance) and |cand| (x_norm after acceptance) do not
		// depend on the evaluation: computed while the pixel warps work
		double step_norm = 0, cand_norm = 0;
		const bool eval_ok = leader_eval(sh, last ? 1 : 2, cand, cand_cost, Hm, g, [&]() {
			double d5[5];
#pragma unroll
			for (int k = 0; k < 5; ++k) d5[k] = x[k] - cand[k];
			step_norm = norm5(d5);
			cand_norm = norm5(cand);
		});
		// ComputeCandidatePointAndEvaluateCost: a failed evaluation is an
		// infinitely bad candidate (trust_region_minimizer.cc:731-747)
		const bool cost_ok = isfinite(cand_cost);
		if (!cost_ok) cand_cost = DBL_MAX;
		{
			// ParameterToleranceReached
			if (step_norm <= P.parameter_tolerance * (x_norm + P.parameter_tolerance))
			{
				S.termination = 1;
				break;
			}
		}
		// FunctionToleranceReached
		if (fabs(x_cost - cand_cost) <= P.function_tolerance * x_cost)
		{
			S.termination = 1;
			break;
		}
		// TrustRegionStepEvaluator::StepQuality
		double rho;
		if (!cost_ok)
			rho = -DBL_MAX;
		else
			rho = sel_max((se_cur - cand_cost) * rcp_fast(model_cost_change),
						  (se_ref - cand_cost) * rcp_fast(se_acc_ref + model_cost_change));
		if (!(rho > P.min_relative_decrease))
		{
			// HandleUnsuccessfulStep
			radius *= inv_decrease;
			inv_decrease *= 0.5;
			iter_cost = cand_cost;
			continue;
		}
		// HandleSuccessfulStep
#pragma unroll
		for (int k = 0; k < 5; ++k) x[k] = cand[k];
		x_norm = cand_norm;
		if (!eval_ok)
		{
			// cost finite, Jacobian not: "Residual and Jacobian evaluation failed"
			S.termination = 2;
			failed = true;
			break;
		}
		x_cost = cand_cost;
		if (!last) scale_system(x, false, Hm, g, scale, Hs, gs, P.gradient_tolerance, grad_max_norm);
		{
			const double t = 2.0 * rho - 1.0;
			radius = sel_min(P.max_radius, radius * rcp_fast(sel_max(1.0 / 3.0, 1.0 - t * t * t)));
		}
		inv_decrease = 0.5;
		reuse_diagonal = false;
		se_cur = cand_cost;
		se_acc_cand += model_cost_change;
		se_acc_ref += model_cost_change;
		if (se_cur < se_min)
		{
			se_min = se_cur;
			se_nonmono = 0;
			se_cand = se_cur;
			se_acc_cand = 0;
		}
		else
		{
			++se_nonmono;
			if (se_cur > se_cand)
			{
				se_cand = se_cur;
				se_acc_cand = 0;
			}
		}
		if (se_nonmono == P.max_consecutive_nonmonotonic_steps)
		{
			se_ref = se_cand;
			se_acc_ref = se_acc_cand;
		}
		step_successful = true;
		iter_cost = x_cost;
		iter_gmn = grad_max_norm;
	}
	if (!failed)
	{
#pragma unroll
		for (int k = 0; k < 5; ++k) params[k] = best[k];
	}
	S.final_cost = final_cost;
}

// --------------------------------------------------------------------------
// Per-feature state carried in registers, identical in every thread of the CTA
// --------------------------------------------------------------------------
struct FeatState
{
	double rx, ry, ix, iy;
	double warp[4];
	double flow;
	long long cur_ts, time_last_update, time_without_update;
	int init, lost, num_of_events, counter, tail, size, n_costs, frame_slot;
	int n_traj;  // Patch::trajectory_.size()
};

__device__ __forceinline__ long long to_i32(double v)
{
	// static_cast<int32_t>(double) as x86-64 cvttsd2si performs it
	if (!(v > -2147483649.0 && v < 2147483648.0)) return (long long)INT_MIN;
	return (long long)(int)v;
}

__device__ __forceinline__ bool in_rect(double rx, double ry, int x, int y)
{
	// cv::Rect2d::contains(Point2i -> Point2d)
	return rx <= x && x < rx + SIDE && ry <= y && y < ry + SIDE;
}

// Job (1): Patch::integrateEvents (ft/src/patch.cpp:65-84).  Polarity
// scatter-accumulate of the window (already staged in shared memory) into the
// 25x25 patch with shared-memory integer atomics; containment is re-tested
// against the CURRENT rect; timestamps pass through int32.
__device__ void integrate_window(FeatShared& sh, FeatState& st)
{
	for (int i = threadIdx.x; i < NPIX; i += NT) sh.nabla[i] = 0;
	__syncthreads();
	for (int k = threadIdx.x; k < st.size; k += NT)
	{
		const ebt_event e = sh.ring[(st.tail + k) & RMASK];
		if (in_rect(st.rx, st.ry, e.x, e.y))
		{
			const int px = (int)(e.x - st.rx), py = (int)(e.y - st.ry);
			atomicAdd(&sh.nabla[py * SIDE + px], (int)e.polarity);
		}
	}
	const long long front = sh.ring[(st.tail + st.size - 1) & RMASK].t_us;
	const long long back = sh.ring[st.tail & RMASK].t_us;
	st.cur_ts = to_i32((double)(front + back) * 0.5);
	st.time_last_update = to_i32((double)back);
	__syncthreads();
}

// Patch::updatePatchRect (ft/src/patch.cpp:49-63): centre = warp^-1 * initPoint.
__device__ __forceinline__ void update_rect(FeatState& st)
{
	const double c = st.warp[0], s = st.warp[1];
	const double mx = st.warp[2] * -1.0, my = st.warp[3] * -1.0;
	const double itx = __dadd_rn(__dmul_rn(c, mx), -__dmul_rn(-s, my));
	const double ity = __dadd_rn(__dmul_rn(-s, mx), __dmul_rn(c, my));
	const double ncx = __dadd_rn(__dadd_rn(__dmul_rn(c, st.ix), __dmul_rn(s, st.iy)), itx);
	const double ncy = __dadd_rn(__dadd_rn(__dmul_rn(-s, st.ix), __dmul_rn(c, st.iy)), ity);
	st.rx = ncx - EBT_PATCH_EXTENT;
	st.ry = ncy - EBT_PATCH_EXTENT;
}

// Optimizer::drawCostMap (ft/src/optimizer.cpp:32-60), run at the end of
// optimize() when OptimizerParams::drawCostMap is set: the solve's functor (old
// rect, normalised patch: still in sh.ctx / sh.m) on the 11x11 grid of integer
// translation offsets around the solved warp.  The pose is rebuilt from its
// angle (Pose2d(pose.log().z(), t): SO2::log = atan2, SO2(theta) = (cos, sin)),
// the flow is the float getFlow(); map(y + 5, x + 5) = cv::norm(residuals, L2)
// = sqrt(sum (u_i / sqrt(N) + m_i)^2), N = 1e-5 + sum u^2, from the two reduced
// sums of a cost-only pixel pass.  121 passes: a diagnostic, off by default.
__device__ __noinline__ void cost_map(FeatShared& sh, double w0, double w1, double w2, double w3, double flow,
									   double* __restrict__ out)
{
	const EvalCtx E = sh.ctx;
	const double theta = atan2(w1, w0);
	double x[5];
	sincos(theta, &x[1], &x[0]);
	x[4] = (double)__double2float_rn(flow);
	double vx, vy;
	sincos(x[4], &vy, &vx);
	for (int k = 0; k < 121; ++k)
	{
		const int ox = k / 11 - 5, oy = k % 11 - 5;
		x[2] = (double)ox + w2;
		x[3] = (double)oy + w3;
		double acc[2];
		eval_pixels<false, NT>(E.grid, E.W, E.H, x, vx, vy, E.tlx, E.tly, sh.m, (int)threadIdx.x, acc, sh.tile, E.r0,
							   E.c0, E.staged != 0);
		block_sum<2>(sh, acc);
		if (threadIdx.x == 0)
		{
			const double N = 1e-5 + acc[0];
			const double s = acc[0] / N + 2.0 * acc[1] / sqrt(N) + E.smm;
			out[(oy + 5) * 11 + (ox + 5)] = sqrt(fmax(s, 0.0));
		}
	}
}

// One Optimizer::optimize (ft/src/optimizer.cpp:62-206): integrate, normalise,
// LM solve, cost filter, state update; cm_out != null: the cost map.
__device__ EBT_NOINLINE void optimize_feature(FeatShared& sh, double* costs, FeatState& st,
								 const FrameTable& ft, const DevParams& P, LmSummary& S,
								 double* cm_out = nullptr)
{
	S.filtered = 0;
	integrate_window(sh, st);
	double a1[1];
	a1[0] = 0;
	for (int i = threadIdx.x; i < NPIX; i += NT)
	{
		const double v = (double)sh.nabla[i];
		a1[0] += v * v;
	}
	block_sum<1>(sh, a1);
	// Patch::getNormalizedIntegratedNabla (patch.cpp:156-159): Mat / s is
	// evaluated by OpenCV as Mat * (1/s)
	const double inv = 1.0 / sqrt(a1[0]);
	a1[0] = 0;
	for (int i = threadIdx.x; i < NPIX; i += NT)
	{
		const double v = (double)sh.nabla[i] * inv;
		sh.m[i] = v;
		a1[0] += v * v;
	}
	block_sum<1>(sh, a1);

	// Stage the template-grid texels the solve will sample: the bounding box of the
	// rect warped by the initial estimate, + the 4x4 tap footprint, centred in a
	// TS x TS tile.  Pixels whose taps leave the tile later (or a box that does not
	// fit: large rotation) read global memory as before.
	int r0 = 0, c0 = 0, staged = 0;
	{
		const double c = st.warp[0], s = st.warp[1], tx = st.warp[2], ty = st.warp[3];
		double lox = DBL_MAX, hix = -DBL_MAX, loy = DBL_MAX, hiy = -DBL_MAX;
#pragma unroll
		for (int k = 0; k < 4; ++k)
		{
			const double px = st.rx + ((k & 1) ? SIDE - 1 : 0), py = st.ry + ((k & 2) ? SIDE - 1 : 0);
			const double wx = c * px - s * py + tx, wy = s * px + c * py + ty;
			lox = wx < lox ? wx : lox;
			hix = wx > hix ? wx : hix;
			loy = wy < loy ? wy : loy;
			hiy = wy > hiy ? wy : hiy;
		}
		if (lox > -1e6 && hix < 1e6 && loy > -1e6 && hiy < 1e6)  // also false for NaN
		{
			const int cl = (int)floor(lox) - 1, ch = (int)floor(hix) + 2;
			const int rl = (int)floor(loy) - 1, rh = (int)floor(hiy) + 2;
			if (ch - cl < TS && rh - rl < TS)
			{
				c0 = cl - (TS - (ch - cl + 1)) / 2;
				r0 = rl - (TS - (rh - rl + 1)) / 2;
				staged = 1;
			}
		}
	}
	if (staged)
	{
		const double2* grid = ft.grid[st.frame_slot];
		const int W = P.width, H = P.height;
		for (int idx = threadIdx.x; idx < TS * TS; idx += NT)
		{
			const int i = idx / TS, j = idx - i * TS;
			const int rr = min(max(r0 + i, 0), H - 1), cc = min(max(c0 + j, 0), W - 1);
			stage_texel(sh.tile, idx, __ldg(grid + rr * W + cc));
		}
	}
	if (threadIdx.x == 0)
	{
		EvalCtx& E = sh.ctx;
		E.grid = ft.grid[st.frame_slot];
		E.W = P.width;
		E.H = P.height;
		E.r0 = r0;
		E.c0 = c0;
		E.staged = staged;
		E.tlx = st.rx;
		E.tly = st.ry;
		E.smm = a1[0];
		E.huber_a = P.huber_a;
	}
	// Patch::getFlow() returns float (patch.h:54)
	double x[5] = {st.warp[0], st.warp[1], st.warp[2], st.warp[3],
				   (double)__double2float_rn(st.flow)};
	if (threadIdx.x < 32)
	{
		WALK_T(wl0);
		lm_solve(sh, P, x, S);  // its first post_request publishes sh.ctx
		WALK_T(wl1);
		WALK_ADD(threadIdx.x == 0, 2, wl1 - wl0);
		if (threadIdx.x == 0)
		{
#pragma unroll
			for (int k = 0; k < 5; ++k) sh.res[k] = x[k];
			sh.res[5] = S.final_cost;
			sh.res_i[0] = S.n_successful;
			sh.res_i[1] = S.n_unsuccessful;
			sh.res_i[2] = S.termination;
			sh.res_i[3] = S.n_jac;
			sh.res_i[4] = S.n_cost;
			sh.req_kind = 0;
		}
		handshake();
	}
	else
		lm_worker(sh);
#pragma unroll
	for (int k = 0; k < 5; ++k) x[k] = sh.res[k];
	S.final_cost = sh.res[5];
	S.n_successful = sh.res_i[0];
	S.n_unsuccessful = sh.res_i[1];
	S.termination = sh.res_i[2];
	S.n_jac = sh.res_i[3];
	S.n_cost = sh.res_i[4];
	__syncthreads();

	// addFinalCost + median-of-5 cost filter (optimizer.cpp:127-157)
	if (threadIdx.x == 0) costs[st.n_costs & (EBT_COST_HISTORY - 1)] = S.final_cost;
	st.n_costs++;
	__syncthreads();
	if (st.n_costs >= 6)
	{
		double v[5];
#pragma unroll
		for (int i = 0; i < 5; ++i) v[i] = costs[(st.n_costs - 6 + i) & (EBT_COST_HISTORY - 1)];
#pragma unroll
		for (int i = 1; i < 5; ++i)
#pragma unroll
			for (int j = i; j > 0; --j)
				if (v[j] < v[j - 1])
				{
					const double t = v[j];
					v[j] = v[j - 1];
					v[j - 1] = t;
				}
		if (v[2] > P.optimizer_threshold)
		{
			st.lost = 1;
			S.filtered = 1;
			return;
		}
	}
	st.flow = wrap_flow(x[4]);
#pragma unroll
	for (int k = 0; k < 4; ++k) st.warp[k] = x[k];
	st.init = 1;
	const double ox = st.rx, oy = st.ry;
	update_rect(st);
	const double h = (SIDE - 1) / 2.;  // Patch::toCorner
	const double dx = (st.rx + h) - (ox + h), dy = (st.ry + h) - (oy + h);
	st.time_without_update = (long long)(1e6 / fmax(1e-1, sqrt(dx * dx + dy * dy)));
	st.n_traj++;  // Patch::addTrajectoryPosition (optimizer.cpp:179-181)
	st.counter = 0;
	if (cm_out) cost_map(sh, st.warp[0], st.warp[1], st.warp[2], st.warp[3], st.flow, cm_out);  // optimizer.cpp:191-197
}

// One pixel of the L1 norm updateNumOfEvents takes: nearest-neighbour sample of the latest
// grid through OpenCV's fixed-point rule, |gx a + gy b| (0 outside the image).
__device__ __forceinline__ double num_events_term(const double2* __restrict__ grid, int W, int H, double c, double s,
												  double tx, double ty, double a, double b, int x0, int y0, int i)
{
	const int yy = i / SIDE;
	const int x = x0 + (i - yy * SIDE), y = y0 + yy;
	const int X0 = __double2int_rn(c * x * 1024) + __double2int_rn(__dadd_rn(__dmul_rn(-s, y), tx) * 1024) + 512;
	const int Y0 = __double2int_rn(s * x * 1024) + __double2int_rn(__dadd_rn(__dmul_rn(c, y), ty) * 1024) + 512;
	const int xi = X0 >> 10, yi = Y0 >> 10;
	if (xi >= 0 && xi < W && yi >= 0 && yi < H)
	{
		const double2 g = __ldg(grid + (size_t)yi * W + xi);
		return fabs(__dadd_rn(__dmul_rn(g.x, a), __dmul_rn(g.y, b)));
	}
	return 0.0;
}

// Job (2b): FeatureDetector::updateNumOfEvents (ft/src/feature_detector.cpp:
// 666-712).  cv::warpAffine(flags = WARP_INVERSE_MAP) is nearest-neighbour with
// 10-bit fixed-point source coordinates; only the cropped rect is sampled, from
// the LATEST frame's gradients.
__device__ EBT_NOINLINE void update_num_of_events(FeatShared& sh, FeatState& st, const FrameTable& ft,
									 const DevParams& P)
{
	const int W = P.width, H = P.height;
	const double cx = (st.rx + (st.rx + SIDE)) * 0.5, cy = (st.ry + (st.ry + SIDE)) * 0.5;
	if (cx <= P.border_margin || cy <= P.border_margin || cx >= W - P.border_margin ||
		cy >= H - P.border_margin)
	{
		st.lost = 1;
		return;
	}
	if (st.rx < 0 || st.ry < 0 || st.rx + SIDE >= W || st.ry + SIDE >= H)
	{
		st.num_of_events = min(max(P.init_num_events, P.min_num_events), P.max_num_events);
		return;
	}
	const double2* grid = ft.grid[ft.latest];
	const double c = st.warp[0], s = st.warp[1], tx = st.warp[2], ty = st.warp[3];
	const int x0 = __double2int_rn(st.rx), y0 = __double2int_rn(st.ry);
	const float flow = __double2float_rn(st.flow);
	// std::cos(float) rounds to float (feature_detector.cpp:702-704)
	const double a = 0.6 * (double)__double2float_rn(cos((double)flow));
	const double b = 0.6 * (double)__double2float_rn(sin((double)flow));
	double acc[1];
	acc[0] = 0;
	for (int i = threadIdx.x; i < NPIX; i += NT) acc[0] += num_events_term(grid, W, H, c, s, tx, ty, a, b, x0, y0, i);
	block_sum<1>(sh, acc);
	// size_t sumPatch = norm; Patch::setNumOfEvents clamps (patch.cpp:208-212)
	const unsigned long long n = (unsigned long long)acc[0];
	st.num_of_events = (int)min(max(n, (unsigned long long)P.min_num_events),
								(unsigned long long)P.max_num_events);
}

// Patch::addTrajectoryPosition after an optimisation (optimizer.cpp:179-181): the
// device keeps the two newest trajectory entries, then — when the products are on —
// Patch::integrateMotionCompensatedEvents (ft/src/patch.cpp:87-130): every window
// event is moved along the newest trajectory segment to the window's mid time,
// rounded half-to-even (static_cast<Point2i>(Point2d)), tested against the moved
// rect and accumulated.  Integer work, global atomics on the feature's 625 ints.
struct McArgs
{
	double rx, ry;
	long long cur_ts;
	int n_traj, size, tail;
};
__device__ __noinline__ void motion_compensated(FeatShared& sh, const McArgs st, DevPatch& gp, int products,
											   int* mc_nablas)
{
	if (threadIdx.x == 0)
	{
		const double px = gp.traj_x[1], py = gp.traj_y[1];
		const long long pt = gp.traj_ts[1];
		const double lx = st.rx + (SIDE - 1) / 2., ly = st.ry + (SIDE - 1) / 2.;
		gp.traj_x[0] = px; gp.traj_y[0] = py; gp.traj_ts[0] = pt;
		gp.traj_x[1] = lx; gp.traj_y[1] = ly; gp.traj_ts[1] = st.cur_ts;
		// midTime = currentTimestamp_ = the newest entry's timestamp
		const long long half = to_i32((double)(st.cur_ts - pt) * 0.5);
		const bool run = products && st.n_traj >= 2 && st.size > 0 && st.cur_ts + half >= st.cur_ts &&
						 pt < st.cur_ts;
		sh.req[0] = run ? 1.0 : 0.0;
		sh.req[1] = lx - px;
		sh.req[2] = ly - py;
		sh.req[3] = (double)(st.cur_ts - pt);
	}
	__syncthreads();
	const bool run = sh.req[0] != 0.0;
	const double dirx = sh.req[1], diry = sh.req[2], t_dif = sh.req[3];
	__syncthreads();
	if (!run) return;
	int* mc = mc_nablas + (size_t)gp.store_slot * NPIX;
	for (int i = threadIdx.x; i < NPIX; i += NT) mc[i] = 0;
	__syncthreads();
	const double t = (double)st.cur_ts;
	for (int k = threadIdx.x; k < st.size; k += NT)
	{
		const ebt_event e = sh.ring[(st.tail + k) & RMASK];
		const double sc = __ddiv_rn(t - (double)e.t_us, t_dif);
		const double cxd = __dadd_rn((double)e.x, __dmul_rn(dirx, sc));
		const double cyd = __dadd_rn((double)e.y, __dmul_rn(diry, sc));
		if (!(fabs(cxd) < 2e9 && fabs(cyd) < 2e9)) continue;
		const int cx = __double2int_rn(cxd), cy = __double2int_rn(cyd);
		if (in_rect(st.rx, st.ry, cx, cy))
		{
			const int px = (int)(cx - st.rx), py = (int)(cy - st.ry);
			atomicAdd(&mc[py * SIDE + px], (int)e.polarity);
		}
	}
	__syncthreads();
}

// Patch::warpImage (ft/src/patch.cpp:132-154): predictedNabla_ = -gx cos(flowDir_)
// - gy sin(flowDir_) of the patch's own template gradients, sampled as
// cv::warpAffine(WARP_INVERSE_MAP) does (nearest neighbour, 10-bit fixed point) on
// Rect(patch_).  Left unchanged when the rect is not strictly inside the image.
struct PredArgs
{
	double rx, ry, c, s, tx, ty, flow;
};
__device__ __noinline__ void predicted_nabla(const PredArgs st, const double2* __restrict__ grid, int W, int H,
											 double* __restrict__ out)
{
	if (st.rx < 0 || st.ry < 0 || st.rx + SIDE >= W || st.ry + SIDE >= H) return;
	const double c = st.c, s = st.s, tx = st.tx, ty = st.ty;
	const int x0 = __double2int_rn(st.rx), y0 = __double2int_rn(st.ry);
	const double a = -cos(st.flow), b = -sin(st.flow);
	for (int i = threadIdx.x; i < NPIX; i += NT)
	{
		const int yy = i / SIDE;
		const int x = x0 + (i - yy * SIDE), y = y0 + yy;
		const int X0 = __double2int_rn(c * x * 1024) +
					   __double2int_rn(__dadd_rn(__dmul_rn(-s, y), tx) * 1024) + 512;
		const int Y0 = __double2int_rn(s * x * 1024) +
					   __double2int_rn(__dadd_rn(__dmul_rn(c, y), ty) * 1024) + 512;
		const int xi = X0 >> 10, yi = Y0 >> 10;
		double v = 0.0;
		if (xi >= 0 && xi < W && yi >= 0 && yi < H)
		{
			const double2 g = __ldg(grid + (size_t)yi * W + xi);
			v = __dadd_rn(__dmul_rn(g.x, a), __dmul_rn(g.y, b));
		}
		out[i] = v;
	}
}

__device__ __forceinline__ void load_state(FeatState& st, const DevPatch& p)
{
	st.rx = p.rx; st.ry = p.ry; st.ix = p.ix; st.iy = p.iy;
#pragma unroll
	for (int k = 0; k < 4; ++k) st.warp[k] = p.warp[k];
	st.flow = p.flow;
	st.cur_ts = p.cur_ts;
	st.time_last_update = p.time_last_update;
	st.time_without_update = p.time_without_update;
	st.init = p.init; st.lost = p.lost; st.num_of_events = p.num_of_events;
	st.counter = p.counter; st.tail = p.win_tail; st.size = p.win_size;
	st.n_costs = p.n_costs; st.frame_slot = p.frame_slot;
	st.n_traj = p.n_traj;
}

__device__ __forceinline__ void store_state(const FeatState& st, DevPatch& p)
{
	p.rx = st.rx; p.ry = st.ry;
#pragma unroll
	for (int k = 0; k < 4; ++k) p.warp[k] = st.warp[k];
	p.flow = st.flow;
	p.cur_ts = st.cur_ts;
	p.time_last_update = st.time_last_update;
	p.time_without_update = st.time_without_update;
	p.init = st.init; p.lost = st.lost; p.num_of_events = st.num_of_events;
	p.counter = st.counter; p.win_tail = st.tail; p.win_size = st.size;
	p.n_costs = st.n_costs;
	p.n_traj = st.n_traj;
}

__device__ void write_record_at(ebt_opt_record* records, unsigned long long slot, unsigned long long cap,
								const FeatState& st, int track_id, const LmSummary& S, long long event_index);

__device__ void write_record(ebt_opt_record* records, unsigned long long* n_records,
							 unsigned long long cap, const FeatState& st, int track_id,
							 const LmSummary& S, long long event_index)
{
	write_record_at(records, atomicAdd(n_records, 1ull), cap, st, track_id, S, event_index);
}

// `slot`: the record's place in the push's buffer, reserved with one atomic (by the caller: the
// multi-stream kernel reserves it BEFORE the solve, so the atomic's round trip hides behind it)
__device__ void write_record_at(ebt_opt_record* records, unsigned long long slot, unsigned long long cap,
								const FeatState& st, int track_id, const LmSummary& S, long long event_index)
{
	if (slot >= cap) return;
	ebt_opt_record r;
	r.track_id = track_id;
	r.n_successful = S.n_successful;
	r.n_unsuccessful = S.n_unsuccessful;
	r.termination = S.termination;
	r.lost = st.lost | (S.filtered ? 2 : 0);
	r.num_of_events = st.num_of_events;
	r.ts = st.cur_ts;
	r.event_index = event_index;
	r.x = st.rx + (SIDE - 1) / 2.;
	r.y = st.ry + (SIDE - 1) / 2.;
	r.final_cost = S.final_cost;
#pragma unroll
	for (int k = 0; k < 4; ++k) r.warp[k] = st.warp[k];
	r.flow_dir = (double)__double2float_rn(st.flow);
	records[slot] = r;
}

// --------------------------------------------------------------------------
// Prefilter: one bit per (feature, event) saying whether the event's pixel lies
// in the feature's *candidate region* = its rect at launch grown by `margin`
// pixels on every side.  While the rect stays within `margin` of where it was,
// every event the rect can contain is a candidate, so the feature CTA only has
// to visit set bits (about 2 % of the stream) instead of every event.
//
// HBM/L2-bound integer work.  Grid: x = tiles of 1024 events, y = groups of
// PF_FEATS features.  A CTA stages the packed (x, y) word of its 1024 events in
// shared memory once; each warp then takes one feature at a time and produces
// the tile's 32 mask words with 32 ballots, stored as one coalesced 128-byte
// row segment mask[feature][tile * 32 + lane].
// --------------------------------------------------------------------------
#define PF_TILE 1024
#define PF_THREADS 256
#define PF_FEATS 128

// integer bounds [x0, x1] x [y0, y1] of the candidate region (a superset)
__device__ __forceinline__ int4 candidate_bounds(double rx, double ry, double margin)
{
	int4 b;
	b.x = (int)floor(rx - margin);
	b.y = (int)ceil(rx + margin + SIDE);
	b.z = (int)floor(ry - margin);
	b.w = (int)ceil(ry + margin + SIDE);
	return b;
}

__device__ __forceinline__ void prefilter_tile(const DevPatch* __restrict__ patches, int n_patches,
											   const ebt_event* __restrict__ events, long long n_events, double margin,
											   unsigned* __restrict__ mask, int words_per_row, unsigned tile, int f0,
											   unsigned* xy, int4* bounds)
{
	const long long tile0 = (long long)tile * PF_TILE;
	const unsigned* ev32 = reinterpret_cast<const unsigned*>(events);
	for (int k = threadIdx.x; k < PF_TILE; k += PF_THREADS)
	{
		const long long e = tile0 + k;
		// word 2 of the 16-byte record = x | y << 16; out-of-range events get a
		// coordinate no region contains
		xy[k] = e < n_events ? __ldg(ev32 + 4 * e + 2) : 0xffffffffu;
	}
	for (int k = threadIdx.x; k < PF_FEATS; k += PF_THREADS)
	{
		const int f = f0 + k;
		int4 b = make_int4(1, 0, 1, 0);  // empty
		if (f < n_patches && !patches[f].lost && !patches[f].foreign) b = candidate_bounds(patches[f].rx, patches[f].ry, margin);
		bounds[k] = b;
	}
	__syncthreads();
	const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
	for (int k = warp; k < PF_FEATS && f0 + k < n_patches; k += PF_THREADS / 32)
	{
		const int4 b = bounds[k];
		unsigned mine = 0;
#pragma unroll 8
		for (int j = 0; j < 32; ++j)
		{
			const unsigned v = xy[j * 32 + lane];
			const int x = (int)(v & 0xffffu), y = (int)(v >> 16);
			const unsigned bal = __ballot_sync(0xffffffffu, x >= b.x && x <= b.y && y >= b.z && y <= b.w);
			if (lane == j) mine = bal;
		}
		mask[(size_t)(f0 + k) * words_per_row + (size_t)tile * 32 + lane] = mine;
	}
}

__global__ void __launch_bounds__(PF_THREADS)
prefilter_kernel(const DevPatch* __restrict__ patches, int n_patches,
				 const ebt_event* __restrict__ events, long long n_events, double margin,
				 unsigned* __restrict__ mask, int words_per_row)
{
	__shared__ unsigned xy[PF_TILE];
	__shared__ int4 bounds[PF_FEATS];
	prefilter_tile(patches, n_patches, events, n_events, margin, mask, words_per_row, blockIdx.x,
				   (int)blockIdx.y * PF_FEATS, xy, bounds);
}

// The same for the streams of one multi-stream push: grid z = stream, x and y sized
// for the largest stream; CTAs beyond a stream's own extent leave at once.
__global__ void __launch_bounds__(PF_THREADS)
prefilter_multi_kernel(const StreamDesc* __restrict__ descs)
{
	__shared__ unsigned xy[PF_TILE];
	__shared__ int4 bounds[PF_FEATS];
	const StreamDesc& D = descs[blockIdx.z];
	if ((long long)blockIdx.x * PF_TILE >= D.n_events || (int)blockIdx.y * PF_FEATS >= D.n_patches) return;
	prefilter_tile(D.patches, D.n_patches, D.events, D.n_events, D.margin, D.mask, D.words_per_row, blockIdx.x,
				   (int)blockIdx.y * PF_FEATS, xy, bounds);
}

// Rebuild one feature's mask row from event `from` on, for a candidate region
// around the rect's current position (the rect left the old region).
__device__ void rebuild_mask_row(unsigned* __restrict__ row, const ebt_event* __restrict__ events,
								 long long n_events, long long from, double rx, double ry, double margin)
{
	const int4 b = candidate_bounds(rx, ry, margin);
	const unsigned* ev32 = reinterpret_cast<const unsigned*>(events);
	const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
	const long long n_words = (n_events + 31) >> 5;
	// four mask words per warp and trip: four independent loads in flight per lane
	for (long long w0 = (from >> 5) + 4 * warp; w0 < n_words; w0 += 4 * NW)
	{
		unsigned v[4];
#pragma unroll
		for (int j = 0; j < 4; ++j)
		{
			const long long e = (w0 + j) * 32 + lane;
			v[j] = e < n_events ? __ldg(ev32 + 4 * e + 2) : 0xffffffffu;  // no region holds 65535
		}
		unsigned mine = 0;
#pragma unroll
		for (int j = 0; j < 4; ++j)
		{
			const int x = (int)(v[j] & 0xffffu), y = (int)(v[j] >> 16);
			const unsigned bal = __ballot_sync(0xffffffffu, x >= b.x && x <= b.y && y >= b.z && y <= b.w);
			if (lane == j) mine = bal;
		}
		if (lane < 4 && w0 + lane < n_words) row[w0 + lane] = mine;
	}
	__syncthreads();  // global writes of this CTA are read back by this CTA only
}

// --------------------------------------------------------------------------
// Stream kernel: FeatureDetector::updatePatches (ft/src/feature_detector.cpp:
// 585-619) for a whole batch of events, one CTA per feature.  Between two
// newImage calls a patch's evolution depends only on its own state and the
// ordered event stream, so each CTA replays the per-event loop for its
// feature.  It walks its prefilter mask row in segments of SEG_WORDS words:
// the set bits of a segment are expanded, in stream order, into a candidate
// list in shared memory; candidates are loaded as 16-byte records (one per
// thread per chunk), tested against the CURRENT rect, compacted in order into
// the shared-memory window ring (ballot + prefix), and the optimisation fires
// at exactly the event where Patch::isReady() (patch.cpp:296-299) first holds;
// the rest of the chunk is then re-tested against the moved rect.  When the
// rect leaves the candidate region the CTA rebuilds its own mask row from the
// next event on and carries on.
// CTA b works on feature order[b] (host-sorted, heaviest first).
// --------------------------------------------------------------------------
#define SEG_WORDS NT

__device__ __forceinline__ int window_need(const FeatState& st, const DevParams& P)
{
	// accepted events still needed before isReady() holds
	return max(P.batch_stride - st.counter, st.num_of_events - st.size);
}

__device__ __forceinline__ void window_grow(FeatState& st, int take)
{
	// Patch::addEvent (patch.cpp:37-47): append, trim to numOfEvents_
	const int grown = st.size + take;
	const int kept = min(grown, st.num_of_events);
	st.tail = (st.tail + grown - kept) & RMASK;
	st.size = kept;
	st.counter += take;
}

// MINCTAS = resident CTAs per SM the register budget is cut for.  The launcher
// matches it to the launch: with at most one (two) feature(s) per SM the build
// with 255 (192) registers runs — every value of the pixel pass and most of the
// LM state in registers, no spill traffic on the serial chain — and the
// EBT_MIN_CTAS build (128 registers) when the features outnumber that.
template <int MINCTAS>
__global__ void __launch_bounds__(NT, MINCTAS)
track_batch_kernel(DevPatch* __restrict__ patches, int n_patches, const int* __restrict__ order,
				   ebt_event* __restrict__ rings, int* __restrict__ nablas,
				   const ebt_event* __restrict__ events, long long n_events,
				   long long event_base, unsigned* mask, int words_per_row, double margin,
				   FrameTable ft, DevParams P,
				   ebt_opt_record* __restrict__ records, unsigned long long* n_records,
				   unsigned long long record_cap)
{
	__shared__ FeatShared sh;
	__shared__ double costs[EBT_COST_HISTORY];
	__shared__ unsigned short cand[SEG_WORDS * 32];
	__shared__ int scan[NW];
	if ((int)blockIdx.x >= n_patches) return;
	const int f = order[blockIdx.x];
	DevPatch& gp = patches[f];
	if (gp.lost || gp.foreign) return;  // uniform over the CTA
	unsigned long long t_enter;
	asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t_enter));
	WALK_T(wk_enter);
	FeatState st;
	load_state(st, gp);
	const int track_id = gp.track_id;
	ebt_event* gring = rings + (size_t)gp.store_slot * EBT_RING;
	int* gnabla = nablas + (size_t)gp.store_slot * NPIX;
	for (int k = threadIdx.x; k < st.size; k += NT) sh.ring[k] = gring[(st.tail + k) & RMASK];
	if (threadIdx.x < EBT_COST_HISTORY) costs[threadIdx.x] = gp.costs[threadIdx.x];
	st.tail = 0;
	__syncthreads();

	const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
	const uint4* ev4 = reinterpret_cast<const uint4*>(events);
	unsigned* row = mask + (size_t)f * words_per_row;
	double srx = st.rx, sry = st.ry;  // where the candidate region is centred
	long long start = 0;              // first stream event not consumed yet
	int n_opt = 0;
	int buf = 0;

	// One Optimizer::optimize at stream event `at`, then the bookkeeping around it.
	// Returns true when the mask row was rebuilt (the caller re-expands).
	auto optimise_at = [&](long long at) -> bool {
		LmSummary S;
		WALK_T(w0);
		optimize_feature(sh, costs, st, ft, P, S,
						 P.cost_maps ? P.cost_maps + (size_t)gp.store_slot * 121 : nullptr);
		WALK_T(w1);
		WALK_ADD(threadIdx.x == 0, 1, w1 - w0);
		WALK_ADD(threadIdx.x == 0, 5, 1);
		if (!S.filtered)
			motion_compensated(sh, McArgs{st.rx, st.ry, st.cur_ts, st.n_traj, st.size, st.tail}, gp, P.products,
							   P.mc_nablas);
		WALK_T(w2);
		update_num_of_events(sh, st, ft, P);
		WALK_T(w3);
		WALK_ADD(threadIdx.x == 0, 6, w3 - w2);
		if (P.products)  // feature_detector.cpp:615
			predicted_nabla(PredArgs{st.rx, st.ry, st.warp[0], st.warp[1], st.warp[2], st.warp[3], st.flow},
							ft.grid[st.frame_slot], P.width, P.height,
							P.pred_nablas + (size_t)gp.store_slot * NPIX);
		if (threadIdx.x == 0) write_record(records, n_records, record_cap, st, track_id, S, event_base + at);
		++n_opt;
		start = at + 1;
		if (!st.lost && (fabs(st.rx - srx) > margin || fabs(st.ry - sry) > margin))
		{
			WALK_T(w4);
			rebuild_mask_row(row, events, n_events, start, st.rx, st.ry, margin);
			WALK_T(w5);
			WALK_ADD(threadIdx.x == 0, 3, w5 - w4);
			WALK_ADD(threadIdx.x == 0, 4, 1);
			srx = st.rx;
			sry = st.ry;
			return true;
		}
		return false;
	};

	// Already ready at entry (newImage shrank numOfEvents_): the test at
	// feature_detector.cpp:611 fires on the very next stream event, accepted or
	// not.  Event 0 is then the only candidate of a first pass.
	bool pre = st.init && n_events > 0 && window_need(st, P) <= 0;

	const long long n_words = (n_events + 31) >> 5;
	long long seg = start >> 5;
	while (!st.lost && seg < n_words)
	{
		// expand the segment's set bits into cand[], in stream order
		WALK_ADD(threadIdx.x == 0, 7, 1);
		const long long wi = seg + threadIdx.x;
		unsigned word = wi < n_words ? row[wi] : 0u;
		{
			const long long first = wi << 5;
			if (first < start) word = (start - first >= 32) ? 0u : (word & (0xffffffffu << (int)(start - first)));
		}
		const int c = __popc(word);
		int incl = c;
#pragma unroll
		for (int o = 1; o < 32; o <<= 1)
		{
			const int up = __shfl_up_sync(0xffffffffu, incl, o);
			if (lane >= o) incl += up;
		}
		if (lane == 31) scan[warp] = incl;
		__syncthreads();
		int off = incl - c, n_c = 0;
#pragma unroll
		for (int w = 0; w < NW; ++w)
		{
			const int v = scan[w];
			off += (w < warp) ? v : 0;
			n_c += v;
		}
		if (pre)
		{
			n_c = 1;
			if (threadIdx.x == 0) cand[0] = 0;
		}
		else
			while (word)
			{
				const int bit = __ffs(word) - 1;
				word &= word - 1;
				cand[off++] = (unsigned short)(threadIdx.x * 32 + bit);
			}
		__syncthreads();

		bool restart = false;
		for (int cbase = 0; cbase < n_c && !restart && !st.lost; cbase += NT)
		{
			const int chunk_len = min(NT, n_c - cbase);
			uint4 raw = make_uint4(0, 0, 0, 0);
			if ((int)threadIdx.x < chunk_len) raw = __ldg(ev4 + ((seg << 5) + cand[cbase + threadIdx.x]));
			const int ex = (int)(raw.z & 0xffffu), ey = (int)(raw.z >> 16);
			int resume = 0;
			while (resume < chunk_len)
			{
				const bool live = (int)threadIdx.x >= resume && (int)threadIdx.x < chunk_len;
				const bool inside = live && in_rect(st.rx, st.ry, ex, ey);
				const unsigned bal = __ballot_sync(0xffffffffu, inside);
				if (lane == 0) sh.bal[buf][warp] = bal;
				__syncthreads();
				unsigned b[NW];
				int before = 0, total = 0;
#pragma unroll
				for (int w = 0; w < NW; ++w)
				{
					b[w] = sh.bal[buf][w];
					const int cw = __popc(b[w]);
					before += (w < warp) ? cw : 0;
					total += cw;
				}
				buf ^= 1;
				const int rank = before + __popc(bal & ((1u << lane) - 1u)) + 1;  // if inside
				const int need = window_need(st, P);  // <= 0 only in the `pre` pass
				int trig = -1, take = total;
				if (st.init && need <= 0)
				{
					trig = resume;
					take = (int)((b[resume >> 5] >> (resume & 31)) & 1u);
				}
				else if (st.init && total >= need)
				{
					take = need;
					int acc = 0;
#pragma unroll
					for (int w = 0; w < NW; ++w)
					{
						const int cw = __popc(b[w]);
						if (trig < 0 && acc + cw >= need) trig = w * 32 + (int)__fns(b[w], 0, need - acc);
						acc += cw;
					}
				}
				if (inside && rank <= take)
					*reinterpret_cast<uint4*>(&sh.ring[(st.tail + st.size + rank - 1) & RMASK]) = raw;
				if (take > 0) window_grow(st, take);
				if (trig < 0) break;
				const long long at = (seg << 5) + cand[cbase + trig];
				if (optimise_at(at))
				{
					restart = true;
					break;
				}
				if (st.lost) break;
				resume = trig + 1;
			}
		}
		__syncthreads();  // cand[] is rewritten by the next expansion
		if (pre)
		{
			pre = false;  // event 0 is consumed; expand the real first segment
			restart = true;
		}
		seg = restart ? (start >> 5) : seg + SEG_WORDS;
	}
	__syncthreads();
	for (int k = threadIdx.x; k < st.size; k += NT) gring[k] = sh.ring[(st.tail + k) & RMASK];
	if (n_opt > 0)
		for (int i = threadIdx.x; i < NPIX; i += NT) gnabla[i] = sh.nabla[i];
	if (threadIdx.x < EBT_COST_HISTORY) gp.costs[threadIdx.x] = costs[threadIdx.x];
	st.tail = 0;
	if (threadIdx.x == 0)
	{
		store_state(st, gp);
		gp.last_n_opt = n_opt;  // load estimate for the next launch's CTA order
		WALK_T(wk_exit);
		WALK_ADD(true, 0, wk_exit - wk_enter);
		unsigned long long t_exit;
		unsigned sm;
		asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t_exit));
		asm volatile("mov.u32 %0, %%smid;" : "=r"(sm));
		gp.last_t0_ns = t_enter;
		gp.last_t1_ns = t_exit;
		gp.last_sm = (int)sm;
	}
}

// --------------------------------------------------------------------------
// Batched kernel: F independent feature-patch problems, one CTA each — one
// Optimizer::optimize (ft/src/optimizer.cpp:62-206) per problem.
// --------------------------------------------------------------------------

// --------------------------------------------------------------------------
// Batched kernel, shared pixel warps: SB_LEADERS leader warps — one feature-patch
// problem each at a time — share ONE set of SB_PIXW pixel warps.  While one
// leader runs its scalar trust-region step, the pixel warps evaluate a sibling's
// request, so they stay busy (in the one-CTA-per-feature kernels a feature's four
// pixel warps idle through its scalar half).  ~12 k registers per feature
// instead of 20 k.  Measured on 65 536 problems (profiles/r1_variants.md): 2 leaders
// + 4 pixel warps at 170 registers, 2 CTAs per SM (4 problems in flight per SM)
// is the fastest shape, +18 % over one CTA per problem; 3 or 4 leaders per pixel
// set fall to 128-144 registers and lose more to spills than they gain.
// The window, integration, normalisation
// and post-solve of a problem run on its leader warp alone; the pixel passes and
// the tile staging are posted as jobs.
//
// Hand-off without barriers: a leader writes its request, then bumps
// slot.req_seq; every pixel warp serves every job (its 32 of the 128 pixel
// threads' share) in the order it discovers them and arrives on the slot's mbarrier; the
// leader waits for that phase (mbarrier.try_wait).  A leader has at most one job
// in flight, so the pixel warps need not agree on an order.
// --------------------------------------------------------------------------
#ifndef SB_LEADERS
#define SB_LEADERS 2
#endif
#ifndef SB_SLEEP
#define SB_SLEEP 0  // ns of back-off per poll; plain spinning measured as fast
#endif
#define SB_BACKOFF() do { if (SB_SLEEP > 0) __nanosleep(SB_SLEEP); } while (0)
#ifndef SB_PIXW
#define SB_PIXW 4
#endif
#ifndef SB_CTAS
#define SB_CTAS 2
#endif
#define SB_THREADS ((SB_LEADERS + SB_PIXW) * 32)
#define SB_RING 304  // >= Patch::maxNumOfEvents_
#ifndef SB_TILE
#define SB_TILE 1
#endif

struct __align__(16) SlotShared
{
	static constexpr int kPixWarps = SB_PIXW;
	static constexpr bool is_slot = true;
	ebt_event ring[SB_RING];  // the problem's window, chronological
#if SB_TILE
	double2 tile[TS * TS];
#else
	double2 tile[1];          // taps through L1
#endif
	double m[NPIX];
	int nabla[NPIX];
	double red[SB_PIXW][20];
	double req[8];
	double tot[20];
	EvalCtx ctx;
	int req_kind;             // 1 cost-only pass, 2 cost + Jacobian pass, 3 stage the tile, 4 leader has finished
	unsigned long long done_bar;  // mbarrier: SB_PIXW pixel-warp arrivals complete a job
	unsigned req_seq;         // jobs posted so far (written by the leader)
	unsigned seq_local;       // leader's own copy of req_seq
	long long t_last;         // EBT_PHASE_CLOCKS only
};

__device__ __forceinline__ unsigned ld_volatile_u32(const unsigned* p) { return *(const volatile unsigned*)p; }

// Completion side of the hand-off: an mbarrier per slot (sm_90+ hardware barrier in shared
// memory).  Each of the SB_PIXW pixel warps arrives once per job; the leader waits for the
// phase of its job with mbarrier.try_wait, which parks the warp in hardware instead of
// spinning on a shared-memory word (the polls took 12 % of the issued instructions of
// the spinning version, profiles/r2_summary.md).  A leader has one job in flight and
// observes every phase, so the phase parity is (jobs posted - 1) & 1.
__device__ __forceinline__ void mbar_init(unsigned long long* bar, unsigned count)
{
	asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"((unsigned)__cvta_generic_to_shared(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_arrive(unsigned long long* bar)
{
	asm volatile("mbarrier.arrive.release.cta.shared::cta.b64 _, [%0];" ::"r"((unsigned)__cvta_generic_to_shared(bar)) : "memory");
}
__device__ __forceinline__ void mbar_wait(unsigned long long* bar, unsigned parity)
{
	asm volatile(
		"{\n"
		".reg .pred P1;\n"
		"EBT_MBAR_WAIT:\n"
		"mbarrier.try_wait.parity.acquire.cta.shared::cta.b64 P1, [%0], %1;\n"
		"@P1 bra EBT_MBAR_DONE;\n"
		"bra EBT_MBAR_WAIT;\n"
		"EBT_MBAR_DONE:\n"
		"}" ::"r"((unsigned)__cvta_generic_to_shared(bar)),
		"r"(parity)
		: "memory");
}

template <class SL>
__device__ __forceinline__ void sb_post(SL& S)
{
	__syncwarp();
	if ((threadIdx.x & 31) == 0)
	{
		__threadfence_block();  // request data before the sequence number
		*(volatile unsigned*)&S.req_seq = ++S.seq_local;
	}
	__syncwarp();
}
template <class SL, class = decltype(SL::is_slot)>
__device__ __forceinline__ void wait_partials(SL& S)
{
	mbar_wait(&S.done_bar, (S.seq_local - 1u) & 1u);
	__syncwarp();
}
template <class SL, class = decltype(SL::is_slot)>
__device__ __forceinline__ void post_request(SL& S, int kind, const double (&x)[5], double vx, double vy)
{
	if ((threadIdx.x & 31) == 0)
	{
#pragma unroll
		for (int k = 0; k < 5; ++k) S.req[k] = x[k];
		S.req[5] = vx;
		S.req[6] = vy;
		S.req_kind = kind;
	}
	sb_post(S);
}

// One job of slot S on this pixel warp (ptid = index among the 128 pixel threads).
template <class SL>
__device__ __noinline__ void sb_serve(SL& S, int kind, int ptid)
{
	const int lane = ptid & 31, pw = ptid >> 5;
	const EvalCtx& E = S.ctx;
	if (kind == 3)
	{
		for (int idx = ptid; idx < TS * TS; idx += SB_PIXW * 32)
		{
			const int i = idx / TS, j = idx - i * TS;
			const int rr = min(max(E.r0 + i, 0), E.H - 1), cc = min(max(E.c0 + j, 0), E.W - 1);
			stage_texel(S.tile, idx, __ldg(E.grid + rr * E.W + cc));
		}
		return;
	}
	double x[5];
#pragma unroll
	for (int k = 0; k < 5; ++k) x[k] = S.req[k];
	const double vx = S.req[5], vy = S.req[6];
	if (kind == 2)
	{
		double acc[20];
		eval_pixels<true, SB_PIXW * 32>(E.grid, E.W, E.H, x, vx, vy, E.tlx, E.tly, S.m, ptid, acc, S.tile, E.r0, E.c0,
										  E.staged != 0);
		Butterfly<20, 16>::run(acc, lane);
		const int slot = butterfly_slot<20>(lane);
		if (slot >= 0) S.red[pw][slot] = acc[0];
	}
	else
	{
		double acc[2];
		eval_pixels<false, SB_PIXW * 32>(E.grid, E.W, E.H, x, vx, vy, E.tlx, E.tly, S.m, ptid, acc, S.tile, E.r0, E.c0,
										   E.staged != 0);
		Butterfly<2, 16>::run(acc, lane);
		const int slot = butterfly_slot<2>(lane);
		if (slot >= 0) S.red[pw][slot] = acc[0];
	}
}

// One Optimizer::optimize of problem `pr` on a leader warp (32 lanes), slot S.
__device__ void sb_optimize(SlotShared& S, const ebt_problem& pr, const ebt_event* __restrict__ events,
							const FrameTable& ft, const DevParams& P, ebt_solution& out, int* __restrict__ nab_out)
{
	const int lane = threadIdx.x & 31;
	FeatState st;
	st.ix = pr.init_x;
	st.iy = pr.init_y;
#pragma unroll
	for (int k = 0; k < 4; ++k) st.warp[k] = pr.warp[k];
	st.flow = pr.flow_dir;
	st.cur_ts = st.time_last_update = 0;
	st.time_without_update = 10000000;
	st.init = 1;
	st.lost = 0;
	st.num_of_events = min(max(pr.n_events, P.min_num_events), P.max_num_events);
	st.size = min(pr.n_events, st.num_of_events);
	st.counter = pr.n_events;
	st.tail = 0;
	st.n_costs = 0;
	st.n_traj = 0;
	st.frame_slot = pr.frame_slot;
	st.rx = st.ry = 0;
	update_rect(st);
	// window given newest first; ring is chronological
	const uint4* ev4 = reinterpret_cast<const uint4*>(events + pr.event_offset);
	for (int k = lane; k < st.size; k += 32) *reinterpret_cast<uint4*>(&S.ring[k]) = __ldg(ev4 + (st.size - 1 - k));
	for (int i = lane; i < NPIX; i += 32) S.nabla[i] = 0;
	__syncwarp();
	// Patch::integrateEvents (patch.cpp:65-84), as integrate_window
	for (int k = lane; k < st.size; k += 32)
	{
		const ebt_event e = S.ring[k];
		if (in_rect(st.rx, st.ry, e.x, e.y))
		{
			const int px = (int)(e.x - st.rx), py = (int)(e.y - st.ry);
			atomicAdd(&S.nabla[py * SIDE + px], (int)e.polarity);
		}
	}
	{
		const long long front = S.ring[st.size - 1].t_us, back = S.ring[0].t_us;
		st.cur_ts = to_i32((double)(front + back) * 0.5);
		st.time_last_update = to_i32((double)back);
	}
	__syncwarp();
	// Patch::getNormalizedIntegratedNabla (patch.cpp:156-159)
	double a = 0;
	for (int i = lane; i < NPIX; i += 32)
	{
		const double v = (double)S.nabla[i];
		a += v * v;
	}
	a = warp_sum(a);
	const double inv = 1.0 / sqrt(a);
	double smm = 0;
	for (int i = lane; i < NPIX; i += 32)
	{
		const double v = (double)S.nabla[i] * inv;
		S.m[i] = v;
		smm += v * v;
	}
	smm = warp_sum(smm);
	// tile origin: as in optimize_feature
	int r0 = 0, c0 = 0, staged = 0;
	{
		const double c = st.warp[0], s = st.warp[1], tx = st.warp[2], ty = st.warp[3];
		double lox = DBL_MAX, hix = -DBL_MAX, loy = DBL_MAX, hiy = -DBL_MAX;
#pragma unroll
		for (int k = 0; k < 4; ++k)
		{
			const double px = st.rx + ((k & 1) ? SIDE - 1 : 0), py = st.ry + ((k & 2) ? SIDE - 1 : 0);
			const double wx = c * px - s * py + tx, wy = s * px + c * py + ty;
			lox = wx < lox ? wx : lox;
			hix = wx > hix ? wx : hix;
			loy = wy < loy ? wy : loy;
			hiy = wy > hiy ? wy : hiy;
		}
		if (SB_TILE && lox > -1e6 && hix < 1e6 && loy > -1e6 && hiy < 1e6)
		{
			const int cl = (int)floor(lox) - 1, ch = (int)floor(hix) + 2;
			const int rl = (int)floor(loy) - 1, rh = (int)floor(hiy) + 2;
			if (ch - cl < TS && rh - rl < TS)
			{
				c0 = cl - (TS - (ch - cl + 1)) / 2;
				r0 = rl - (TS - (rh - rl + 1)) / 2;
				staged = 1;
			}
		}
	}
	if (lane == 0)
	{
		EvalCtx& E = S.ctx;
		E.grid = ft.grid[st.frame_slot];
		E.W = P.width;
		E.H = P.height;
		E.r0 = r0;
		E.c0 = c0;
		E.staged = staged;
		E.tlx = st.rx;
		E.tly = st.ry;
		E.smm = smm;
		E.huber_a = P.huber_a;
		S.req_kind = 3;
	}
	if (staged)
	{
		sb_post(S);
		wait_partials(S);
	}
	double x[5] = {st.warp[0], st.warp[1], st.warp[2], st.warp[3], (double)__double2float_rn(st.flow)};
	LmSummary L;
	L.filtered = 0;
	lm_solve(S, P, x, L);
	// post-solve (optimizer.cpp:127-189); one final cost => the cost filter cannot fire
	st.flow = wrap_flow(x[4]);
#pragma unroll
	for (int k = 0; k < 4; ++k) st.warp[k] = x[k];
	update_rect(st);
	if (nab_out)
		for (int i = lane; i < NPIX; i += 32) nab_out[i] = S.nabla[i];
	if (lane == 0)
	{
		ebt_solution s;
#pragma unroll
		for (int k = 0; k < 4; ++k) s.warp[k] = st.warp[k];
		s.flow_dir = (double)__double2float_rn(st.flow);
		s.final_cost = L.final_cost;
		s.x = st.rx + (SIDE - 1) / 2.;
		s.y = st.ry + (SIDE - 1) / 2.;
		s.n_successful = L.n_successful;
		s.n_unsuccessful = L.n_unsuccessful;
		s.termination = L.termination;
		s.n_passes = L.n_jac * 100 + L.n_cost;
		s.ts = st.cur_ts;
		out = s;
	}
	__syncwarp();
}

#ifdef SB_MAXNREG
__global__ void __maxnreg__(SB_MAXNREG)
#else
__global__ void __launch_bounds__(SB_THREADS, SB_CTAS)
#endif
optimize_batch_shared_kernel(const ebt_problem* __restrict__ problems, int n_problems,
							 const ebt_event* __restrict__ events, FrameTable ft, DevParams P,
							 ebt_solution* __restrict__ solutions, int* __restrict__ nablas)
{
	extern __shared__ __align__(16) unsigned char sb_raw[];
	SlotShared* slots = reinterpret_cast<SlotShared*>(sb_raw);
	const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
	if (threadIdx.x < SB_LEADERS)
	{
		slots[threadIdx.x].req_seq = 0;
		mbar_init(&slots[threadIdx.x].done_bar, SB_PIXW);
		slots[threadIdx.x].seq_local = 0;
		slots[threadIdx.x].t_last = 0;
	}
	__syncthreads();
	if (warp < SB_LEADERS)
	{
		SlotShared& S = slots[warp];
		for (int b = blockIdx.x * SB_LEADERS + warp; b < n_problems; b += gridDim.x * SB_LEADERS)
			sb_optimize(S, problems[b], events, ft, P, solutions[b], nablas ? nablas + (size_t)b * NPIX : nullptr);
		if (lane == 0) S.req_kind = 4;
		sb_post(S);
		return;
	}
	const int ptid = threadIdx.x - SB_LEADERS * 32;
	unsigned served[SB_LEADERS];
#pragma unroll
	for (int l = 0; l < SB_LEADERS; ++l) served[l] = 0;
	int exited = 0;
	while (exited < SB_LEADERS)
	{
		bool idle = true;
#pragma unroll
		for (int l = 0; l < SB_LEADERS; ++l)
		{
			SlotShared& S = slots[l];
			const unsigned seq = __shfl_sync(0xffffffffu, ld_volatile_u32(&S.req_seq), 0);
			if (seq == served[l]) continue;
			served[l] = seq;
			idle = false;
			__threadfence_block();  // the request behind the sequence number
			const int kind = *(const volatile int*)&S.req_kind;
			if (kind == 4)
			{
				++exited;
				continue;
			}
			sb_serve(S, kind, ptid);
			__threadfence_block();
			__syncwarp();
			if (lane == 0) mbar_arrive(&S.done_bar);
		}
		if (idle) SB_BACKOFF();
	}
}


// --------------------------------------------------------------------------
// Stream kernel, shared pixel warps, many streams: FeatureDetector::updatePatches
// (ft/src/feature_detector.cpp:585-619) for the event batches of S independent
// streams in ONE launch.  A work item is one feature of one stream; persistent
// CTAs of two leader warps + four pixel warps pull work items, heaviest first,
// from a queue.  A leader warp owns one feature at a time and keeps only what is
// serial on itself: the candidate walk over the feature's prefilter row and the
// scalar trust-region chain.  Everything that is parallel over pixels or window
// events is posted to the CTA's pixel warps as a job (hand-off by sequence
// numbers, as in optimize_batch_shared_kernel):
//   5  Patch::integrateEvents: the 625 pixels are dealt out to the pixel warps
//      (pixel i belongs to warp i mod 4), every warp scans the whole window and
//      accumulates the events that fall on its own pixels, then sums their squares
//      — no ordering between the pixel warps is needed; the same job stages the
//      template-grid tile of the solve (the leader then normalises the patch itself)
//   1/2 cost / cost + Jacobian pixel pass of the LM solve
//   6  the 625 nearest-neighbour samples of FeatureDetector::updateNumOfEvents
// While one leader walks, or runs its scalar step, the pixel warps serve the other.
// Candidate walk: 128 mask words (4096 stream events) per segment, one 16-byte
// load per lane, the words of the next two segments already in flight.
// Shape (measured, profiles/r2_variants.md): two leaders per four pixel warps at 168
// registers, two CTAs per SM.  Three leaders per pixel set — uniformly at 144 registers,
// or with setmaxnreg moving registers from the leader warpgroup (88-112) to the pixel
// warpgroup (144-168) — were 25-40 % slower: the leaders' scalar trust-region chain is
// the critical path and it is the part that cannot give registers away.
// --------------------------------------------------------------------------
#define MQ_SEG_WORDS 128
struct __align__(16) StreamSlot
{
	static constexpr int kPixWarps = SB_PIXW;
	static constexpr bool is_slot = true;
	ebt_event ring[EBT_RING];
	double2 tile[TS * TS];
	double m[NPIX];
	int nabla[NPIX];
	double red[SB_PIXW][20];
	double req[8];
	double tot[20];
	double costs[EBT_COST_HISTORY];
	EvalCtx ctx;
	unsigned short cand[MQ_SEG_WORDS * 32];
	const double2* job_grid;  // job 6: the latest frame's grid
	// of the feature's stream, read once per feature instead of once per optimisation
	const double2* const* grids;
	ebt_opt_record* records;
	unsigned long long* n_records;
	unsigned long long record_cap;
	long long event_base;
	int latest;
	int pad_;
	// the two newest entries of Patch::trajectory_ (DevPatch::traj_*), written back at the end
	double traj_x[2], traj_y[2];
	long long traj_ts[2];
	// the feature's state between optimisations and where its walk stands in the batch: in shared
	// memory so that nothing of it has to survive the solve in registers (see the leader loop)
	FeatState st;
	struct
	{
		double srx, sry;  // where the candidate region of the mask row is centred
		double margin;
		long long start;  // first stream event not consumed yet
		long long n_events;
		const ebt_event* events;
		unsigned* row;
		unsigned long long t_enter;
		int wpr, n_opt, track_id, win_tail0;
	} walk;
	int job_i[4];             // integer job arguments (5: window tail, size; 6: x0, y0)
	int redi[SB_PIXW];        // job 5: per-warp sum of squares of the integer patch
	unsigned long long done_bar;  // mbarrier: SB_PIXW pixel-warp arrivals complete a job
	int req_kind;
	unsigned req_seq, seq_local;
	long long t_last;
};

// Jobs 3, 5 and 6 of slot S on this pixel warp (ptid = index among the 128 pixel threads).
__device__ __noinline__ void mq_serve(StreamSlot& S, int kind, int ptid)
{
	const int lane = ptid & 31, pw = ptid >> 5;
	const EvalCtx& E = S.ctx;
	if (kind == 5)
	{
		// Patch::integrateEvents (patch.cpp:65-84) on this warp's pixels
		const double rx = S.req[0], ry = S.req[1];
		const int tail = S.job_i[0], size = S.job_i[1];
		for (int i = pw + SB_PIXW * lane; i < NPIX; i += SB_PIXW * 32) S.nabla[i] = 0;
		__syncwarp();
		for (int k = lane; k < size; k += 32)
		{
			const ebt_event e = S.ring[(tail + k) & RMASK];
			if (in_rect(rx, ry, e.x, e.y))
			{
				const int idx = (int)(e.y - ry) * SIDE + (int)(e.x - rx);
				if ((idx & (SB_PIXW - 1)) == pw) atomicAdd(&S.nabla[idx], (int)e.polarity);
			}
		}
		__syncwarp();
		int ss = 0;
		for (int i = pw + SB_PIXW * lane; i < NPIX; i += SB_PIXW * 32)
		{
			const int v = S.nabla[i];
			ss += v * v;
		}
#pragma unroll
		for (int o = 16; o > 0; o >>= 1) ss += __shfl_xor_sync(0xffffffffu, ss, o);
		if (lane == 0) S.redi[pw] = ss;
		// ... and, in the same job, the template-grid tile of the solve (independent of the integration)
		if (E.staged)
			for (int idx = ptid; idx < TS * TS; idx += SB_PIXW * 32)
			{
				const int i = idx / TS, j = idx - i * TS;
				const int rr = min(max(E.r0 + i, 0), E.H - 1), cc = min(max(E.c0 + j, 0), E.W - 1);
				stage_texel(S.tile, idx, __ldg(E.grid + rr * E.W + cc));
			}
		return;
	}
	// kind 6: the L1 norm FeatureDetector::updateNumOfEvents takes (feature_detector.cpp:695-705)
	const double c = S.req[0], s = S.req[1], tx = S.req[2], ty = S.req[3], a = S.req[4], b = S.req[5];
	const int x0 = S.job_i[0], y0 = S.job_i[1];
	const double2* grid = S.job_grid;
	double acc = 0;
	for (int i = ptid; i < NPIX; i += SB_PIXW * 32) acc += num_events_term(grid, E.W, E.H, c, s, tx, ty, a, b, x0, y0, i);
	acc = warp_sum(acc);
	if (lane == 0) S.red[pw][0] = acc;
}

// FeatureDetector::updateNumOfEvents on a leader warp, the pixel sum as job 6
__device__ void mq_update_num_of_events(StreamSlot& S, FeatState& st, const double2* grid_latest, const DevParams& P)
{
	const int lane = threadIdx.x & 31;
	const int W = P.width, H = P.height;
	const double cx = (st.rx + (st.rx + SIDE)) * 0.5, cy = (st.ry + (st.ry + SIDE)) * 0.5;
	if (cx <= P.border_margin || cy <= P.border_margin || cx >= W - P.border_margin || cy >= H - P.border_margin)
	{
		st.lost = 1;
		return;
	}
	if (st.rx < 0 || st.ry < 0 || st.rx + SIDE >= W || st.ry + SIDE >= H)
	{
		st.num_of_events = min(max(P.init_num_events, P.min_num_events), P.max_num_events);
		return;
	}
	const float flow = __double2float_rn(st.flow);
	if (lane == 0)
	{
		S.req[0] = st.warp[0];
		S.req[1] = st.warp[1];
		S.req[2] = st.warp[2];
		S.req[3] = st.warp[3];
		// std::cos(float) rounds to float (feature_detector.cpp:702-704)
		S.req[4] = 0.6 * (double)__double2float_rn(cos((double)flow));
		S.req[5] = 0.6 * (double)__double2float_rn(sin((double)flow));
		S.job_i[0] = __double2int_rn(st.rx);
		S.job_i[1] = __double2int_rn(st.ry);
		S.job_grid = grid_latest;
		S.ctx.W = W;
		S.ctx.H = H;
		S.req_kind = 6;
	}
	sb_post(S);
	wait_partials(S);
	const double acc = red_total(S, 0);
	const unsigned long long n = (unsigned long long)acc;
	st.num_of_events = (int)min(max(n, (unsigned long long)P.min_num_events), (unsigned long long)P.max_num_events);
}

// One Optimizer::optimize on a leader warp (as optimize_feature, cost map excluded)
__device__ void mq_optimize(StreamSlot& S, FeatState& st, const double2* grid, const DevParams& P, LmSummary& L)
{
	const int lane = threadIdx.x & 31;
	L.filtered = 0;
	// tile origin: the bounding box of the rect warped by the initial estimate, + the 4x4 tap footprint
	int r0 = 0, c0 = 0, staged = 0;
	{
		const double c = st.warp[0], s = st.warp[1], tx = st.warp[2], ty = st.warp[3];
		double lox = DBL_MAX, hix = -DBL_MAX, loy = DBL_MAX, hiy = -DBL_MAX;
#pragma unroll
		for (int k = 0; k < 4; ++k)
		{
			const double px = st.rx + ((k & 1) ? SIDE - 1 : 0), py = st.ry + ((k & 2) ? SIDE - 1 : 0);
			const double wx = c * px - s * py + tx, wy = s * px + c * py + ty;
			lox = wx < lox ? wx : lox;
			hix = wx > hix ? wx : hix;
			loy = wy < loy ? wy : loy;
			hiy = wy > hiy ? wy : hiy;
		}
		if (lox > -1e6 && hix < 1e6 && loy > -1e6 && hiy < 1e6)
		{
			const int cl = (int)floor(lox) - 1, ch = (int)floor(hix) + 2;
			const int rl = (int)floor(loy) - 1, rh = (int)floor(hiy) + 2;
			if (ch - cl < TS && rh - rl < TS)
			{
				c0 = cl - (TS - (ch - cl + 1)) / 2;
				r0 = rl - (TS - (rh - rl + 1)) / 2;
				staged = 1;
			}
		}
	}
	// job 5: the pixel warps integrate the window and stage the tile
	if (lane == 0)
	{
		EvalCtx& E = S.ctx;
		E.grid = grid;
		E.W = P.width;
		E.H = P.height;
		E.r0 = r0;
		E.c0 = c0;
		E.staged = staged;
		E.tlx = st.rx;
		E.tly = st.ry;
		E.huber_a = P.huber_a;
		S.req[0] = st.rx;
		S.req[1] = st.ry;
		S.job_i[0] = st.tail;
		S.job_i[1] = st.size;
		S.req_kind = 5;
	}
	sb_post(S);
	// meanwhile: timestamps
	{
		const long long front = S.ring[(st.tail + st.size - 1) & RMASK].t_us;
		const long long back = S.ring[st.tail & RMASK].t_us;
		st.cur_ts = to_i32((double)(front + back) * 0.5);
		st.time_last_update = to_i32((double)back);
	}
	wait_partials(S);
	// the sum of squares of an integer patch is an exact integer in any order
	long long ssq = 0;
#pragma unroll
	for (int w = 0; w < SB_PIXW; ++w) ssq += (long long)S.redi[w];
	// Patch::getNormalizedIntegratedNabla (patch.cpp:156-159): Mat * (1/s), on the leader itself —
	// 20 trips are cheaper than another job's round trip behind the sibling's pixel pass
	const double inv = 1.0 / sqrt((double)ssq);
	{
		double a = 0;
		for (int i = lane; i < NPIX; i += 32)
		{
			const double v = (double)S.nabla[i] * inv;
			S.m[i] = v;
			a += v * v;
		}
		a = warp_sum(a);
		if (lane == 0) S.ctx.smm = a;
		__syncwarp();
	}
	double x[5] = {st.warp[0], st.warp[1], st.warp[2], st.warp[3], (double)__double2float_rn(st.flow)};
	WALK_T(wl0);
	lm_solve(S, P, x, L);
	WALK_T(wl1);
	WALK_ADD(lane == 0, 2, wl1 - wl0);
	// addFinalCost + median-of-5 cost filter (optimizer.cpp:127-157)
	if (lane == 0) S.costs[st.n_costs & (EBT_COST_HISTORY - 1)] = L.final_cost;
	st.n_costs++;
	__syncwarp();
	if (st.n_costs >= 6)
	{
		double v[5];
#pragma unroll
		for (int i = 0; i < 5; ++i) v[i] = S.costs[(st.n_costs - 6 + i) & (EBT_COST_HISTORY - 1)];
#pragma unroll
		for (int i = 1; i < 5; ++i)
#pragma unroll
			for (int j = i; j > 0; --j)
				if (v[j] < v[j - 1])
				{
					const double t = v[j];
					v[j] = v[j - 1];
					v[j - 1] = t;
				}
		if (v[2] > P.optimizer_threshold)
		{
			st.lost = 1;
			L.filtered = 1;
			return;
		}
	}
	st.flow = wrap_flow(x[4]);
#pragma unroll
	for (int k = 0; k < 4; ++k) st.warp[k] = x[k];
	st.init = 1;
	const double ox = st.rx, oy = st.ry;
	update_rect(st);
	const double h = (SIDE - 1) / 2.;
	const double dx = (st.rx + h) - (ox + h), dy = (st.ry + h) - (oy + h);
	st.time_without_update = (long long)(1e6 / fmax(1e-1, sqrt(dx * dx + dy * dy)));
	st.n_traj++;
	st.counter = 0;
}

// this lane's four words of the 128-word mask segment that starts at word `seg` (a multiple of 4)
__device__ __forceinline__ uint4 mq_load_segment(const unsigned* row, long long seg, int words_per_row, int lane)
{
	const long long w = seg + 4 * lane;
	if (w < words_per_row) return __ldcg(reinterpret_cast<const uint4*>(row + w));  // rows are whole 128-byte tiles
	return make_uint4(0, 0, 0, 0);
}

__global__ void __launch_bounds__(SB_THREADS, SB_CTAS)
track_multi_kernel(const StreamDesc* __restrict__ descs, const WorkItem* __restrict__ work, int n_work, DevParams P,
				   int* __restrict__ queue)
{
	extern __shared__ __align__(16) unsigned char ss_raw[];
	StreamSlot* slots = reinterpret_cast<StreamSlot*>(ss_raw);
	const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
	if (threadIdx.x < SB_LEADERS)
	{
		slots[threadIdx.x].req_seq = 0;
		mbar_init(&slots[threadIdx.x].done_bar, SB_PIXW);
		slots[threadIdx.x].seq_local = 0;
		slots[threadIdx.x].t_last = 0;
	}
	__syncthreads();
	if (warp >= SB_LEADERS)
	{
		const int ptid = threadIdx.x - SB_LEADERS * 32;
		unsigned served[SB_LEADERS];
#pragma unroll
		for (int l = 0; l < SB_LEADERS; ++l) served[l] = 0;
		int exited = 0;
		while (exited < SB_LEADERS)
		{
			bool idle = true;
#pragma unroll
			for (int l = 0; l < SB_LEADERS; ++l)
			{
				StreamSlot& S = slots[l];
				const unsigned seq = __shfl_sync(0xffffffffu, ld_volatile_u32(&S.req_seq), 0);
				if (seq == served[l]) continue;
				served[l] = seq;
				idle = false;
				__threadfence_block();
				const int kind = *(const volatile int*)&S.req_kind;
				if (kind == 4)
				{
					++exited;
					continue;
				}
				if (kind <= 2)
					sb_serve(S, kind, ptid);
				else
					mq_serve(S, kind, ptid);
				__threadfence_block();
				__syncwarp();
				if (lane == 0) mbar_arrive(&S.done_bar);
			}
			if (idle) SB_BACKOFF();
		}
		return;
	}
	// ---- leader warp `warp`.  First work item: CTA c pairs the c-th heaviest with the c-th
	// lightest (work[] is heaviest first), so two long chains do not share pixel warps;
	// afterwards the leaders pull the items in between from the queue, heaviest first.
	//
	// What the walk needs of the feature's state between optimisations (rect, window cursor,
	// counters, where it stands in the batch) lives in the slot's shared memory (S.st, S.walk), not in
	// registers: values that would have to survive the inlined solve in registers get spilled to local
	// memory and come back through an L1 that the streaming loads keep evicting (measured: 3.4 k cycles
	// per ballot round).  Discipline: every lane reads at the top of a round, lane 0 writes behind the
	// round's ballot, __syncwarp() separates the write from the next round's reads.
	StreamSlot& S = slots[warp];
	const int n_ctas = (int)gridDim.x;
#if SB_LEADERS == 2
	int oi = warp == 0 ? (int)blockIdx.x : n_work - 1 - (int)blockIdx.x;
	if (warp == 1 && oi < n_ctas) oi = n_work;  // that one belongs to a first leader
	for (; oi < n_work;
		 oi = __shfl_sync(0xffffffffu, lane == 0 ? atomicAdd(queue, 1) : 0, 0), oi = oi < n_work - n_ctas ? oi : n_work)
#else
	// other shapes (measurement builds): leader 0 starts on the CTA's own item, everyone else on the queue
	int oi = warp == 0 ? (int)blockIdx.x : __shfl_sync(0xffffffffu, lane == 0 ? atomicAdd(queue, 1) : 0, 0);
	for (; oi < n_work; oi = __shfl_sync(0xffffffffu, lane == 0 ? atomicAdd(queue, 1) : 0, 0))
#endif
	{
		const WorkItem item = work[oi];
		const StreamDesc& D = descs[item.stream];
		DevPatch& gp = D.patches[item.feature];
		if (gp.lost || gp.foreign) continue;
		WALK_T(wk_enter);
		if (lane == 0)
		{
			asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(S.walk.t_enter));
			load_state(S.st, gp);
			S.walk.win_tail0 = S.st.tail;
			S.st.tail = 0;
			S.walk.srx = S.st.rx;
			S.walk.sry = S.st.ry;
			S.walk.start = 0;
			S.walk.n_opt = 0;
			S.walk.track_id = gp.track_id;
			S.walk.margin = D.margin;
			S.walk.events = D.events;
			S.walk.n_events = D.n_events;
			S.walk.row = D.mask + (size_t)item.feature * D.words_per_row;
			S.walk.wpr = D.words_per_row;
			S.grids = D.grids;
			S.records = D.records;
			S.n_records = D.n_records;
			S.record_cap = D.record_cap;
			S.event_base = D.event_base;
			S.latest = D.latest;
		}
		if (lane < EBT_COST_HISTORY) S.costs[lane] = gp.costs[lane];
		if (lane < 2)
		{
			S.traj_x[lane] = gp.traj_x[lane];
			S.traj_y[lane] = gp.traj_y[lane];
			S.traj_ts[lane] = gp.traj_ts[lane];
		}
		__syncwarp();
		{
			const ebt_event* gring = D.rings + (size_t)gp.store_slot * EBT_RING;
			const int tail0 = S.walk.win_tail0, size0 = S.st.size;
			for (int k = lane; k < size0; k += 32) S.ring[k] = gring[(tail0 + k) & RMASK];
		}
		__syncwarp();
		const long long n_events = S.walk.n_events;
		const long long n_words = (n_events + 31) >> 5;
		bool pre = S.st.init && n_events > 0 && window_need(S.st, P) <= 0;
		long long seg = 0;
		uint4 cur = mq_load_segment(S.walk.row, seg, S.walk.wpr, lane);
		uint4 nxt = mq_load_segment(S.walk.row, seg + MQ_SEG_WORDS, S.walk.wpr, lane);
		while (!S.st.lost && seg < n_words)
		{
			WALK_T(ws0);
			const unsigned* row = S.walk.row;
			const int wpr = S.walk.wpr;
			const long long start = S.walk.start;
			// the words of the next two segments travel while this one is walked
			const uint4 nxt2 = mq_load_segment(row, seg + 2 * MQ_SEG_WORDS, wpr, lane);
			unsigned w4[4] = {cur.x, cur.y, cur.z, cur.w};
			int c = 0;
#pragma unroll
			for (int j = 0; j < 4; ++j)
			{
				const long long first = (seg + 4 * lane + j) << 5;
				if (first < start) w4[j] = (start - first >= 32) ? 0u : (w4[j] & (0xffffffffu << (int)(start - first)));
				c += __popc(w4[j]);
			}
			int incl = c;
#pragma unroll
			for (int o = 1; o < 32; o <<= 1)
			{
				const int up = __shfl_up_sync(0xffffffffu, incl, o);
				if (lane >= o) incl += up;
			}
			int off = incl - c;
			int n_c = __shfl_sync(0xffffffffu, incl, 31);
			if (pre)
			{
				n_c = 1;
				if (lane == 0) S.cand[0] = 0;
			}
			else
			{
#pragma unroll
				for (int j = 0; j < 4; ++j)
				{
					unsigned word = w4[j];
					while (word)
					{
						const int bit = __ffs(word) - 1;
						word &= word - 1;
						S.cand[off++] = (unsigned short)((4 * lane + j) * 32 + bit);
					}
				}
			}
			__syncwarp();
			WALK_T(ws1);
			WALK_ADD(lane == 0, 3, ws1 - ws0);  // diagnostic builds: slot 3 = mask wait + expansion
			bool restart = false, lost = false;
			const uint4* ev4 = reinterpret_cast<const uint4*>(S.walk.events);
			// candidate records: chunk c + 1 is loaded before chunk c is walked
			uint4 raw_next = make_uint4(0, 0, 0, 0);
			if (lane < n_c) raw_next = __ldg(ev4 + ((seg << 5) + S.cand[lane]));
			for (int cbase = 0; cbase < n_c && !restart && !lost; cbase += 32)
			{
				const int chunk_len = min(32, n_c - cbase);
				const uint4 raw = raw_next;
				if (cbase + 32 + lane < n_c) raw_next = __ldg(ev4 + ((seg << 5) + S.cand[cbase + 32 + lane]));
				const int ex = (int)(raw.z & 0xffffu), ey = (int)(raw.z >> 16);
				int resume = 0;
				while (resume < chunk_len)
				{
					// this round's view of the window state (lane 0 rewrites it behind the ballot)
					const double rx = S.st.rx, ry = S.st.ry;
					const int w_init = S.st.init, w_counter = S.st.counter, w_size = S.st.size, w_tail = S.st.tail,
							  w_noe = S.st.num_of_events;
					const bool live = lane >= resume && lane < chunk_len;
					const bool inside = live && in_rect(rx, ry, ex, ey);
					const unsigned bal = __ballot_sync(0xffffffffu, inside);
					const int total = __popc(bal);
					const int rank = __popc(bal & ((1u << lane) - 1u)) + 1;  // if inside
					const int need = max(P.batch_stride - w_counter, w_noe - w_size);  // window_need
					int trig = -1, take = total;
					if (w_init && need <= 0)
					{
						trig = resume;
						take = (int)((bal >> resume) & 1u);
					}
					else if (w_init && total >= need)
					{
						take = need;
						trig = (int)__fns(bal, 0, need);
					}
					if (inside && rank <= take)
						*reinterpret_cast<uint4*>(&S.ring[(w_tail + w_size + rank - 1) & RMASK]) = raw;
					if (take > 0 && lane == 0)
					{
						// Patch::addEvent (patch.cpp:37-47): append, trim to numOfEvents_ (window_grow)
						const int grown = w_size + take;
						const int kept = min(grown, w_noe);
						S.st.tail = (w_tail + grown - kept) & RMASK;
						S.st.size = kept;
						S.st.counter = w_counter + take;
					}
					__syncwarp();
					if (trig < 0) break;
					const long long at = (seg << 5) + S.cand[cbase + trig];
					// ---- one Optimizer::optimize at stream event `at` and the bookkeeping around it, on a
					// register copy of the state that lives only inside this block
					{
						FeatState st = S.st;
						LmSummary L;
						WALK_T(w0);
						// the record's slot is reserved now: the atomic's round trip hides behind the solve
						unsigned long long rec_slot = 0;
						if (lane == 0) rec_slot = atomicAdd(S.n_records, 1ull);
						mq_optimize(S, st, S.grids[st.frame_slot], P, L);
						WALK_T(w1);
						WALK_ADD(lane == 0, 1, w1 - w0);
						WALK_ADD(lane == 0, 5, 1);
						if (!L.filtered && lane == 0)
						{
							// Patch::addTrajectoryPosition: the two newest trajectory entries
							S.traj_x[0] = S.traj_x[1]; S.traj_y[0] = S.traj_y[1]; S.traj_ts[0] = S.traj_ts[1];
							S.traj_x[1] = st.rx + (SIDE - 1) / 2.; S.traj_y[1] = st.ry + (SIDE - 1) / 2.; S.traj_ts[1] = st.cur_ts;
						}
						WALK_T(w2);
						mq_update_num_of_events(S, st, S.grids[S.latest], P);
						WALK_T(w3);
						WALK_ADD(lane == 0, 6, w3 - w2);
						lost = st.lost != 0;
						const bool drifted = !lost && (fabs(st.rx - S.walk.srx) > S.walk.margin || fabs(st.ry - S.walk.sry) > S.walk.margin);
						__syncwarp();  // every lane has taken its copy and read the walk state
						if (lane == 0)
						{
							write_record_at(S.records, rec_slot, S.record_cap, st, S.walk.track_id, L, S.event_base + at);
							S.st = st;
							S.walk.n_opt++;
							S.walk.start = at + 1;
							if (drifted)
							{
								S.walk.srx = st.rx;
								S.walk.sry = st.ry;
							}
						}
						__syncwarp();
						if (drifted)
						{
							// rebuild the mask row from the next event on around the rect's current position
							const int4 bb = candidate_bounds(st.rx, st.ry, S.walk.margin);
							const unsigned* ev32 = reinterpret_cast<const unsigned*>(S.walk.events);
							unsigned* wrow = S.walk.row;
							// eight mask words per trip: eight independent loads in flight per lane
							for (long long w0 = ((at + 1) >> 5); w0 < n_words; w0 += 8)
							{
								unsigned v[8];
#pragma unroll
								for (int j = 0; j < 8; ++j)
								{
									const long long e = (w0 + j) * 32 + lane;
									v[j] = e < n_events ? __ldg(ev32 + 4 * e + 2) : 0xffffffffu;  // no region holds 65535
								}
								unsigned mine = 0;
#pragma unroll
								for (int j = 0; j < 8; ++j)
								{
									const int x = (int)(v[j] & 0xffffu), y = (int)(v[j] >> 16);
									const unsigned b2 =
										__ballot_sync(0xffffffffu, x >= bb.x && x <= bb.y && y >= bb.z && y <= bb.w);
									if (lane == j) mine = b2;
								}
								if (lane < 8 && w0 + lane < n_words) __stcg(wrow + w0 + lane, mine);
							}
							__syncwarp();
							restart = true;
						}
					}
					if (restart || lost) break;
					resume = trig + 1;
				}
			}
			__syncwarp();
			WALK_T(ws2);
			WALK_ADD(lane == 0, 4, ws2 - ws1);  // diagnostic builds: slot 4 = chunk phase incl. optimisations
			if (pre)
			{
				pre = false;
				restart = true;
			}
			if (restart)
			{
				seg = (S.walk.start >> 5) & ~3LL;
				cur = mq_load_segment(S.walk.row, seg, S.walk.wpr, lane);
				nxt = mq_load_segment(S.walk.row, seg + MQ_SEG_WORDS, S.walk.wpr, lane);
			}
			else
			{
				seg += MQ_SEG_WORDS;
				cur = nxt;
				nxt = nxt2;
			}
		}
		__syncwarp();
		{
			ebt_event* gring = D.rings + (size_t)gp.store_slot * EBT_RING;
			int* gnabla = D.nablas + (size_t)gp.store_slot * NPIX;
			const int tail = S.st.tail, size = S.st.size, n_opt = S.walk.n_opt;
			for (int k = lane; k < size; k += 32) gring[k] = S.ring[(tail + k) & RMASK];
			if (n_opt > 0)
				for (int i = lane; i < NPIX; i += 32) gnabla[i] = S.nabla[i];
			if (lane < EBT_COST_HISTORY) gp.costs[lane] = S.costs[lane];
			if (lane < 2 && n_opt > 0)
			{
				gp.traj_x[lane] = S.traj_x[lane];
				gp.traj_y[lane] = S.traj_y[lane];
				gp.traj_ts[lane] = S.traj_ts[lane];
			}
			__syncwarp();
			if (lane == 0)
			{
				S.st.tail = 0;
				store_state(S.st, gp);
				gp.last_n_opt = n_opt;
				WALK_T(wk_exit);
				WALK_ADD(true, 0, wk_exit - wk_enter);
				unsigned long long t_exit;
				unsigned sm;
				asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t_exit));
				asm volatile("mov.u32 %0, %%smid;" : "=r"(sm));
				gp.last_t0_ns = S.walk.t_enter;
				gp.last_t1_ns = t_exit;
				gp.last_sm = (int)sm;
			}
		}
		__syncwarp();
	}
	if (lane == 0) S.req_kind = 4;
	sb_post(S);
}


template <int MINCTAS>  // register budget matched to the launch, as for track_batch_kernel
__global__ void __launch_bounds__(NT, MINCTAS)
optimize_batch_kernel(const ebt_problem* __restrict__ problems, int n_problems,
					  const ebt_event* __restrict__ events, FrameTable ft, DevParams P,
					  ebt_solution* __restrict__ solutions, int* __restrict__ nablas)
{
	__shared__ FeatShared sh;
	__shared__ double costs[EBT_COST_HISTORY];
	const int b = blockIdx.x;
	if (b >= n_problems) return;
	const ebt_problem pr = problems[b];
	FeatState st;
	st.ix = pr.init_x;
	st.iy = pr.init_y;
#pragma unroll
	for (int k = 0; k < 4; ++k) st.warp[k] = pr.warp[k];
	st.flow = pr.flow_dir;
	st.cur_ts = st.time_last_update = 0;
	st.time_without_update = 10000000;
	st.init = 1;
	st.lost = 0;
	st.num_of_events = min(max(pr.n_events, P.min_num_events), P.max_num_events);
	st.size = min(pr.n_events, st.num_of_events);
	st.counter = pr.n_events;
	st.tail = 0;
	st.n_costs = 0;
	st.n_traj = 0;
	st.frame_slot = pr.frame_slot;
	st.rx = st.ry = 0;
	update_rect(st);
	// window given newest first; ring is chronological
	const uint4* ev4 = reinterpret_cast<const uint4*>(events + pr.event_offset);
	for (int k = threadIdx.x; k < st.size; k += NT)
		*reinterpret_cast<uint4*>(&sh.ring[k]) = __ldg(ev4 + (st.size - 1 - k));
	__syncthreads();
	LmSummary S;
	optimize_feature(sh, costs, st, ft, P, S, P.cost_maps ? P.cost_maps + (size_t)b * 121 : nullptr);
	if (nablas)
		for (int i = threadIdx.x; i < NPIX; i += NT) nablas[(size_t)b * NPIX + i] = sh.nabla[i];
	if (threadIdx.x == 0)
	{
		ebt_solution s;
#pragma unroll
		for (int k = 0; k < 4; ++k) s.warp[k] = st.warp[k];
		s.flow_dir = (double)__double2float_rn(st.flow);
		s.final_cost = S.final_cost;
		s.x = st.rx + (SIDE - 1) / 2.;
		s.y = st.ry + (SIDE - 1) / 2.;
		s.n_successful = S.n_successful;
		s.n_unsuccessful = S.n_unsuccessful;
		s.termination = S.termination;
		s.n_passes = S.n_jac * 100 + S.n_cost;
		s.ts = st.cur_ts;
		solutions[b] = s;
	}
}

// FeatureDetector::newImage loop "updateNumOfEvents(patch)" for every patch
// (ft/src/feature_detector.cpp:499-502), one CTA per patch.
__global__ void __launch_bounds__(NT)
num_events_kernel(DevPatch* __restrict__ patches, int n_patches, FrameTable ft, DevParams P)
{
	__shared__ FeatShared sh;
	const int f = blockIdx.x;
	if (f >= n_patches) return;
	FeatState st;
	load_state(st, patches[f]);
	update_num_of_events(sh, st, ft, P);
	if (P.products == 2)  // feature_detector.cpp:504-510
		predicted_nabla(PredArgs{st.rx, st.ry, st.warp[0], st.warp[1], st.warp[2], st.warp[3], st.flow},
						ft.grid[st.frame_slot], P.width, P.height,
						P.pred_nablas + (size_t)patches[f].store_slot * NPIX);
	if (threadIdx.x == 0)
	{
		patches[f].lost = st.lost;
		patches[f].num_of_events = st.num_of_events;
	}
}

// Job (2a): FeatureDetector::getLogImage + getGradients (ft/src/
// feature_detector.cpp:714-732) fused with Optimizer::setGrad's interleave
// (ft/src/optimizer.cpp:15-30): L = lut[I] (= log(I/255 + 0.1)/8, 256-entry
// table uploaded by the host), Sobel 3x3 with BORDER_REFLECT_101, row pass
// then column pass, written as (gx, gy) pairs.
__global__ void grad_kernel(const unsigned char* __restrict__ gray, const double* __restrict__ lut,
							int W, int H, double2* __restrict__ grid)
{
	__shared__ double slut[256];
	for (int i = threadIdx.y * blockDim.x + threadIdx.x; i < 256; i += blockDim.x * blockDim.y)
		slut[i] = lut[i];
	__syncthreads();
	const int x = blockIdx.x * blockDim.x + threadIdx.x;
	const int y = blockIdx.y * blockDim.y + threadIdx.y;
	if (x >= W || y >= H) return;
	const int xm = x - 1 < 0 ? 1 : x - 1, xp = x + 1 >= W ? 2 * W - 2 - (x + 1) : x + 1;
	const int ym = y - 1 < 0 ? 1 : y - 1, yp = y + 1 >= H ? 2 * H - 2 - (y + 1) : y + 1;
	double dif[3], smo[3];
	const int ys[3] = {ym, y, yp};
#pragma unroll
	for (int r = 0; r < 3; ++r)
	{
		const unsigned char* row = gray + (size_t)ys[r] * W;
		const double a = slut[row[xm]], b = slut[row[x]], c = slut[row[xp]];
		dif[r] = c - a;
		smo[r] = a + b * 2 + c;
	}
	double2 g;
	g.x = dif[0] + dif[1] * 2 + dif[2];
	g.y = smo[2] - smo[0];
	grid[(size_t)y * W + x] = g;
}

// Roofline denominator for the LM kernels: dependent-chain FMA throughput of
// the non-tensor FP64 / FP32 pipes (8 independent chains per thread).
template <typename T>
__global__ void fma_peak_kernel(T* out, int iters, T a, T b)
{
	T v[8];
#pragma unroll
	for (int k = 0; k < 8; ++k) v[k] = (T)(threadIdx.x + k);
	for (int i = 0; i < iters; ++i)
	{
#pragma unroll
		for (int k = 0; k < 8; ++k) v[k] = v[k] * a + b;
	}
	T s = 0;
#pragma unroll
	for (int k = 0; k < 8; ++k) s += v[k];
	out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}
}  // namespace

// multiprocessors of the current device (every device of a box is the same part)
static int device_sm_count()
{
	static int sm_count = 0;
	if (!sm_count)
	{
		int dev = 0;
		cudaGetDevice(&dev);
		cudaDeviceGetAttribute(&sm_count, cudaDevAttrMultiProcessorCount, dev);
	}
	return sm_count;
}

cudaError_t ebt_launch_fma_peak(int fp64, void* out, int blocks, int threads, int iters,
								cudaStream_t stream)
{
	if (fp64)
		fma_peak_kernel<double><<<blocks, threads, 0, stream>>>((double*)out, iters, 1.0000001, 1e-9);
	else
		fma_peak_kernel<float><<<blocks, threads, 0, stream>>>((float*)out, iters, 1.0000001f, 1e-9f);
	return cudaGetLastError();
}

// --------------------------------------------------------------------------
// Launchers (called from ebt_api.cu)
// --------------------------------------------------------------------------
cudaError_t ebt_launch_track_batch(DevPatch* patches, int n_patches, const int* order,
								   ebt_event* rings, int* nablas, const ebt_event* events,
								   long long n_events, long long event_base, unsigned* mask,
								   int words_per_row, double margin, const FrameTable& ft,
								   const DevParams& P, ebt_opt_record* records,
								   unsigned long long* n_records, unsigned long long record_cap,
								   cudaStream_t stream)
{
	if (n_patches <= 0 || n_events <= 0) return cudaSuccess;
	const dim3 pf_grid((unsigned)((n_events + PF_TILE - 1) / PF_TILE),
					   (unsigned)((n_patches + PF_FEATS - 1) / PF_FEATS));
	prefilter_kernel<<<pf_grid, PF_THREADS, 0, stream>>>(patches, n_patches, events, n_events, margin,
														  mask, words_per_row);
	const int sm_count = device_sm_count();
#define EBT_TRACK_LAUNCH(M)                                                                                 \
	track_batch_kernel<M><<<n_patches, NT, 0, stream>>>(patches, n_patches, order, rings, nablas, events,  \
														n_events, event_base, mask, words_per_row, margin, \
														ft, P, records, n_records, record_cap)
	static const int forced = getenv("EBT_STREAM_BUILD") ? atoi(getenv("EBT_STREAM_BUILD")) : 0;  // tuning only
	const int build = forced ? forced : (n_patches <= sm_count ? 1 : (n_patches <= 2 * sm_count ? 2 : 3));
	if (build == 1)
		EBT_TRACK_LAUNCH(1);
	else if (build == 2 && EBT_MIN_CTAS > 2)
		EBT_TRACK_LAUNCH(2);
	else
		EBT_TRACK_LAUNCH(EBT_MIN_CTAS);
#undef EBT_TRACK_LAUNCH
	return cudaGetLastError();
}

// Prefilter + stream kernel for the event batches of n_streams streams in one launch each.
// descs: device array of n_streams descriptors; work: device array of n_work (stream, feature)
// items, heaviest first; max_events / max_patches: the largest stream's (prefilter grid).
cudaError_t ebt_launch_track_multi(const StreamDesc* descs, int n_streams, const WorkItem* work, int n_work,
								   long long max_events, int max_patches, const DevParams& P, int* queue,
								   cudaStream_t stream)
{
	if (n_streams <= 0 || n_work <= 0 || max_events <= 0) return cudaSuccess;
	const dim3 pf_grid((unsigned)((max_events + PF_TILE - 1) / PF_TILE), (unsigned)((max_patches + PF_FEATS - 1) / PF_FEATS),
					   (unsigned)n_streams);
	prefilter_multi_kernel<<<pf_grid, PF_THREADS, 0, stream>>>(descs);
	const size_t smem = SB_LEADERS * sizeof(StreamSlot);
	int dev = 0;
	cudaGetDevice(&dev);
	static unsigned long long attr_devices = 0;
	if (!(attr_devices >> dev & 1ull))
	{
		cudaError_t e = cudaFuncSetAttribute(track_multi_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
		if (e != cudaSuccess) return e;
		attr_devices |= 1ull << dev;
	}
	const int ctas = min((n_work + SB_LEADERS - 1) / SB_LEADERS, SB_CTAS * device_sm_count());
	cudaError_t e = cudaMemcpyAsync(queue, &ctas, sizeof(int), cudaMemcpyHostToDevice, stream);  // pageable: staged by the driver
	if (e != cudaSuccess) return e;
	track_multi_kernel<<<ctas, SB_THREADS, smem, stream>>>(descs, work, n_work, P, queue);
	return cudaGetLastError();
}

// mask words per feature row for a push of n_events (whole prefilter tiles)
int ebt_mask_words_per_row(long long n_events)
{
	return (int)(((n_events + PF_TILE - 1) / PF_TILE) * (PF_TILE / 32));
}

cudaError_t ebt_launch_optimize_batch(const ebt_problem* problems, int n_problems,
									  const ebt_event* events, const FrameTable& ft,
									  const DevParams& P, ebt_solution* solutions, int* nablas,
									  cudaStream_t stream)
{
	if (n_problems <= 0) return cudaSuccess;
	const int sm_count = device_sm_count();
	// many problems: the shared-pixel-warp kernel (6 problems in flight per SM); the cost map
	// needs the whole CTA per problem and stays on the one-CTA-per-problem kernel
	static const int shared_on = getenv("EBT_BATCH_SHARED") ? atoi(getenv("EBT_BATCH_SHARED")) : 1;
	if (shared_on && !P.cost_maps && n_problems > 4 * sm_count)
	{
		static bool attr_set = false;
		const size_t smem = SB_LEADERS * sizeof(SlotShared);
		if (!attr_set)
		{
			cudaError_t e = cudaFuncSetAttribute(optimize_batch_shared_kernel,
												 cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
			if (e != cudaSuccess) return e;
			attr_set = true;
		}
		const int ctas = min((n_problems + SB_LEADERS - 1) / SB_LEADERS, SB_CTAS * sm_count);
		optimize_batch_shared_kernel<<<ctas, SB_THREADS, smem, stream>>>(problems, n_problems, events, ft, P,
																		 solutions, nablas);
		return cudaGetLastError();
	}
	if (n_problems <= sm_count)
		optimize_batch_kernel<1><<<n_problems, NT, 0, stream>>>(problems, n_problems, events, ft, P, solutions, nablas);
	else if (n_problems <= 2 * sm_count && EBT_MIN_CTAS > 2)
		optimize_batch_kernel<2><<<n_problems, NT, 0, stream>>>(problems, n_problems, events, ft, P, solutions, nablas);
	else
		optimize_batch_kernel<EBT_MIN_CTAS><<<n_problems, NT, 0, stream>>>(problems, n_problems, events, ft, P,
																		solutions, nablas);
	return cudaGetLastError();
}

cudaError_t ebt_launch_num_events(DevPatch* patches, int n_patches, const FrameTable& ft,
								  const DevParams& P, cudaStream_t stream)
{
	if (n_patches <= 0) return cudaSuccess;
	num_events_kernel<<<n_patches, NT, 0, stream>>>(patches, n_patches, ft, P);
	return cudaGetLastError();
}

cudaError_t ebt_launch_grad(const unsigned char* gray, const double* lut, int W, int H,
							double2* grid, cudaStream_t stream)
{
	const dim3 blk(32, 8);
	const dim3 grd((W + blk.x - 1) / blk.x, (H + blk.y - 1) / blk.y);
	grad_kernel<<<grd, blk, 0, stream>>>(gray, lut, W, H, grid);
	return cudaGetLastError();
}

cudaError_t ebt_launch_phase_clocks(unsigned long long* out8, cudaStream_t stream)
{
	cudaError_t e = cudaStreamSynchronize(stream);
	if (e != cudaSuccess) return e;
	e = cudaMemcpyFromSymbol(out8, g_phase_clocks, sizeof(unsigned long long) * 8);
	if (e != cudaSuccess) return e;
	unsigned long long zero[8] = {0, 0, 0, 0, 0, 0, 0, 0};
	return cudaMemcpyToSymbol(g_phase_clocks, zero, sizeof(zero));
}
