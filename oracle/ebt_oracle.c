/* ebt_oracle.c — TEST INFRASTRUCTURE ONLY.
 *
 * Plain-C, single-threaded, FP64 restatement of the reference's event-based
 * feature-tracker hot path.  Only tests/,
 * __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs
 * may load it; the product (CUDA) path never does.
 *
 * Parity status: PINNED.  Checked in tests/test_oracle.py against
 *   (1) the reference's own known-answer test Patch.integrateEventsTest
 *       (implementation/feature_tracker/test/patch_test.cpp:35-60),
 *   (2) outputs of the reference itself (oracle/_ref/libebt_ref.so = the
 *       unmodified patch.cpp/optimizer.cpp + vendored Ceres 2.0, built by
 *       oracle/Makefile) on seeded inputs, frozen under tests/golden/.
 *
 * Every function cites the reference file:line it follows ("ft/" =
 * implementation/feature_tracker/, "ceres/" = thirdparty/ceres-solver/,
 * "sophus/" = thirdparty/Sophus/sophus/).
 */
#define _GNU_SOURCE
#include <float.h>
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

#include "../include/ebt_types.h"

#define NPIX EBT_PATCH_PIXELS
#define SIDE EBT_PATCH_SIDE

/* ------------------------------------------------------------------------ */
/* Gradient grid (template frame)                                            */
/* ------------------------------------------------------------------------ */
typedef struct
{
	int w, h;
	double* g; /* interleaved (gx, gy), row stride w: Optimizer::setGrad,
				  ft/src/optimizer.cpp:15-30 */
	int64_t ts;
	int users; /* Optimizer::used_ */
	int alive;
} oframe;

static int refl101(int i, int n)
{
	if (i < 0) return -i;
	if (i >= n) return 2 * n - 2 - i;
	return i;
}

/* FeatureDetector::getLogImage + getGradients, ft/src/feature_detector.cpp:
 * 714-732: L = log(I*(1/255) + 0.1), Sobel3x3(L/8), BORDER_REFLECT_101;
 * separable, row pass then column pass. */
static void log_sobel(const uint8_t* gray, int w, int h, double* g)
{
	double lut[256];
	size_t n = (size_t)w * h;
	double* L = (double*)malloc(n * sizeof(double));
	double* dif = (double*)malloc(n * sizeof(double));
	double* smo = (double*)malloc(n * sizeof(double));
	for (int i = 0; i < 256; ++i) lut[i] = log(i * (1.0 / 255.0) + 10e-2) * (1.0 / 8);
	for (size_t i = 0; i < n; ++i) L[i] = lut[gray[i]];
	for (int y = 0; y < h; ++y)
		for (int x = 0; x < w; ++x)
		{
			const double a = L[y * w + refl101(x - 1, w)];
			const double b = L[y * w + x];
			const double c = L[y * w + refl101(x + 1, w)];
			dif[y * w + x] = c - a;
			smo[y * w + x] = a + b * 2 + c;
		}
	for (int y = 0; y < h; ++y)
		for (int x = 0; x < w; ++x)
		{
			const int y0 = refl101(y - 1, h), y2 = refl101(y + 1, h);
			g[2 * (y * w + x)] = dif[y0 * w + x] + dif[y * w + x] * 2 + dif[y2 * w + x];
			g[2 * (y * w + x) + 1] = smo[y2 * w + x] - smo[y0 * w + x];
		}
	free(L);
	free(dif);
	free(smo);
}

/* ceres::Grid2D<double,2>::GetValue, ceres/include/ceres/cubic_interpolation.h:
 * 403-421 (clamp to edge). */
static void grid_get(const oframe* f, int r, int c, double* v)
{
	const int ri = r < 0 ? 0 : (r > f->h - 1 ? f->h - 1 : r);
	const int ci = c < 0 ? 0 : (c > f->w - 1 ? f->w - 1 : c);
	v[0] = f->g[2 * (ri * f->w + ci)];
	v[1] = f->g[2 * (ri * f->w + ci) + 1];
}

/* ceres::CubicHermiteSpline<2>, cubic_interpolation.h:64-90. */
static void hermite2(const double* p0, const double* p1, const double* p2,
					 const double* p3, double x, double* f, double* dfdx)
{
	for (int k = 0; k < 2; ++k)
	{
		const double a = 0.5 * (-p0[k] + 3.0 * p1[k] - 3.0 * p2[k] + p3[k]);
		const double b = 0.5 * (2.0 * p0[k] - 5.0 * p1[k] + 4.0 * p2[k] - p3[k]);
		const double c = 0.5 * (-p0[k] + p2[k]);
		const double d = p1[k];
		if (f) f[k] = d + x * (c + x * (b + x * a));
		if (dfdx) dfdx[k] = c + x * (2.0 * b + 3.0 * a * x);
	}
}

/* ceres::BiCubicInterpolator::Evaluate, cubic_interpolation.h:264-325. */
static void bicubic(const oframe* fr, double r, double c, double* f, double* dfdr,
					double* dfdc)
{
	const int row = (int)floor(r);
	const int col = (int)floor(c);
	double p[4][2], fv[4][2], dc[4][2];
	for (int i = 0; i < 4; ++i)
	{
		for (int j = 0; j < 4; ++j) grid_get(fr, row - 1 + i, col - 1 + j, p[j]);
		hermite2(p[0], p[1], p[2], p[3], c - col, fv[i], dc[i]);
	}
	hermite2(fv[0], fv[1], fv[2], fv[3], r - row, f, dfdr);
	if (dfdc) hermite2(dc[0], dc[1], dc[2], dc[3], r - row, dfdc, NULL);
}

/* ------------------------------------------------------------------------ */
/* Cost functor + analytic Jacobian + Huber block correction                 */
/* ------------------------------------------------------------------------ */
typedef struct
{
	const oframe* frame;
	const double* m; /* normalised integrated patch, ft/src/patch.cpp:156-159 */
	double tlx, tly; /* rect top-left at solve start */
	double huber_a;  /* HuberLoss(a): a = 0.3, b = a*a */
} oproblem;

static int all_finite(const double* v, int n)
{
	for (int i = 0; i < n; ++i)
		if (!isfinite(v[i])) return 0;
	return 1;
}

/* OptimizerCostFunctor::operator() + warp(), ft/include/feature_tracker/
 * optimizer_cost.h:29-90, with the analytic Jacobian of SURVEY.md 8a
 * (verified there against Ceres autodiff), the SE2 local chain rule
 * (Dx_this_mul_exp_x_at_0, sophus/se2.hpp:123-142) and the Huber correction
 * of ceres/internal/ceres/residual_block.cc:134-175 / corrector.cc:41-110 /
 * loss_function.cc:48-62.
 *
 * x = [cos, sin, tx, ty, phi].  Outputs: *cost = rho(s)/2; r (625, corrected)
 * and J (625x4 local: upsilon_x, upsilon_y, theta, phi; corrected) when
 * non-NULL.  Returns 0 if any value is non-finite (Ceres: evaluation failed,
 * residual_block.cc:111-131). */
static int evaluate(const oproblem* P, const double* x, double* cost, double* r,
					double* J)
{
	const double cs = x[0], sn = x[1], tx = x[2], ty = x[3];
	const double vx = cos(x[4]), vy = sin(x[4]);
	const int W = P->frame->w, H = P->frame->h;
	double u[NPIX];
	double du[NPIX][4];
	double N = 1e-5;
	double A[4] = {0, 0, 0, 0};
	double rr[NPIX];
	for (int y = 0; y < SIDE; ++y)
		for (int xx = 0; xx < SIDE; ++xx)
		{
			const int i = xx + SIDE * y;
			const double px = xx + P->tlx, py = y + P->tly;
			const double wx = cs * px + -sn * py + tx;
			const double wy = sn * px + cs * py + ty;
			u[i] = 0;
			du[i][0] = du[i][1] = du[i][2] = du[i][3] = 0;
			if (wx >= W || wy >= H || wx < 0.0 || wy < 0.0) continue;
			double g[2], gr[2], gc[2];
			bicubic(P->frame, wy, wx, g, J ? gr : NULL, J ? gc : NULL);
			u[i] = g[0] * vx + g[1] * vy;
			N += u[i] * u[i];
			if (J)
			{
				const double ux = gc[0] * vx + gc[1] * vy; /* du/dWx */
				const double uy = gr[0] * vx + gr[1] * vy; /* du/dWy */
				/* dW/d(upsilon) = R ; dW/d(theta) = R * (-py, px) */
				du[i][0] = ux * cs + uy * sn;
				du[i][1] = -ux * sn + uy * cs;
				du[i][2] = ux * (-cs * py - sn * px) + uy * (-sn * py + cs * px);
				du[i][3] = -g[0] * vy + g[1] * vx;
				for (int q = 0; q < 4; ++q) A[q] += u[i] * du[i][q];
			}
		}
	const double sq = sqrt(N);
	double s = 0;
	for (int i = 0; i < NPIX; ++i)
	{
		rr[i] = u[i] / sq + P->m[i];
		s += rr[i] * rr[i];
	}
	if (!all_finite(rr, NPIX)) return 0;
	/* HuberLoss::Evaluate */
	double rho0, rho1;
	const double b = P->huber_a * P->huber_a;
	if (s > b)
	{
		const double rt = sqrt(s);
		rho0 = 2.0 * P->huber_a * rt - b;
		rho1 = P->huber_a / rt;
		if (rho1 < DBL_MIN) rho1 = DBL_MIN;
	}
	else
	{
		rho0 = s;
		rho1 = 1.0;
	}
	*cost = 0.5 * rho0;
	/* Corrector with rho'' <= 0: scale residuals and Jacobian by sqrt(rho') */
	const double scale = sqrt(rho1);
	if (J)
	{
		const double n32 = N * sq;
		for (int i = 0; i < NPIX; ++i)
			for (int q = 0; q < 4; ++q)
				J[i * 4 + q] = scale * (du[i][q] / sq - u[i] * A[q] / n32);
		if (!all_finite(J, NPIX * 4)) return 0;
	}
	if (r)
		for (int i = 0; i < NPIX; ++i) r[i] = scale * rr[i];
	return 1;
}

/* ------------------------------------------------------------------------ */
/* Parameter update                                                          */
/* ------------------------------------------------------------------------ */
/* LocalParameterizationSE2::Plus = T * SE2::exp(delta), ft/include/
 * feature_tracker/local_parameterization_se2.hpp:20-29; SE2::exp sophus/
 * se2.hpp:560-582; SO2 product + renormalisation sophus/so2.hpp:214-239 and
 * the normalising SO2(real, imag) constructor so2.hpp:388-391.  The flow angle
 * block has no parameterisation: phi + delta[3]. */
static void plus5(const double* x, const double* d, double* out)
{
	const double theta = d[2];
	double ec = cos(theta), es = sin(theta);
	{
		const double len = sqrt(ec * ec + es * es); /* SO2(cos,sin) normalises */
		ec /= len;
		es /= len;
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
		sbt = es / theta;
		omc = (1. - ec) / theta;
	}
	const double ex = sbt * d[0] - omc * d[1];
	const double ey = omc * d[0] + sbt * d[1];
	double rc = x[0] * ec - x[1] * es;
	double rs = x[0] * es + x[1] * ec;
	const double n2 = rc * rc + rs * rs;
	if (n2 != 1.0)
	{
		const double sc = 2.0 / (1.0 + n2);
		rc *= sc;
		rs *= sc;
	}
	{
		const double len = sqrt(rc * rc + rs * rs); /* SO2(real, imag) ctor */
		rc /= len;
		rs /= len;
	}
	out[0] = rc;
	out[1] = rs;
	out[2] = x[2] + (x[0] * ex - x[1] * ey);
	out[3] = x[3] + (x[1] * ex + x[0] * ey);
	out[4] = x[4] + d[3];
}

/* ------------------------------------------------------------------------ */
/* Dense QR least squares: DenseQRSolver::SolveUsingEigen, ceres/internal/   */
/* ceres/dense_qr_solver.cc:120-165  ->  min || [J; diag(D)] y - [r; 0] ||   */
/* by Householder reflections (Eigen householderQr().solve()).               */
/* ------------------------------------------------------------------------ */
static void qr_solve4(const double* J, const double* D, const double* r, double* y)
{
	enum { M = NPIX + 4 };
	static double A[M][4];
	static double b[M];
	for (int i = 0; i < NPIX; ++i)
	{
		for (int q = 0; q < 4; ++q) A[i][q] = J[i * 4 + q];
		b[i] = r[i];
	}
	for (int k = 0; k < 4; ++k)
	{
		for (int q = 0; q < 4; ++q) A[NPIX + k][q] = (q == k) ? D[k] : 0.0;
		b[NPIX + k] = 0;
	}
	for (int k = 0; k < 4; ++k)
	{
		double tail = 0;
		for (int i = k + 1; i < M; ++i) tail += A[i][k] * A[i][k];
		const double c0 = A[k][k];
		double beta, tau;
		if (tail <= DBL_MIN)
		{
			tau = 0;
			beta = c0;
			for (int i = k + 1; i < M; ++i) A[i][k] = 0;
		}
		else
		{
			beta = sqrt(c0 * c0 + tail);
			if (c0 >= 0) beta = -beta;
			for (int i = k + 1; i < M; ++i) A[i][k] /= (c0 - beta);
			tau = (beta - c0) / beta;
		}
		A[k][k] = beta;
		if (tau != 0)
		{
			for (int q = k + 1; q < 4; ++q)
			{
				double w = A[k][q];
				for (int i = k + 1; i < M; ++i) w += A[i][k] * A[i][q];
				w *= tau;
				A[k][q] -= w;
				for (int i = k + 1; i < M; ++i) A[i][q] -= A[i][k] * w;
			}
			double w = b[k];
			for (int i = k + 1; i < M; ++i) w += A[i][k] * b[i];
			w *= tau;
			b[k] -= w;
			for (int i = k + 1; i < M; ++i) b[i] -= A[i][k] * w;
		}
	}
	for (int k = 3; k >= 0; --k)
	{
		double v = b[k];
		for (int q = k + 1; q < 4; ++q) v -= A[k][q] * y[q];
		y[k] = v / A[k][k];
	}
}

/* ------------------------------------------------------------------------ */
/* Trust-region Levenberg-Marquardt, Ceres 2.0 defaults + the reference's    */
/* options (ft/src/optimizer.cpp:102-110): DENSE_QR, non-monotonic steps,    */
/* max 10 iterations.                                                        */
/* ------------------------------------------------------------------------ */
typedef struct
{
	double final_cost;
	int n_successful;   /* accepted steps, iteration 0 not counted */
	int n_unsuccessful; /* rejected + invalid */
	int termination;    /* 0 NO_CONVERGENCE 1 CONVERGENCE 2 FAILURE */
	int n_jac_evals, n_cost_evals;
} osummary;

static void col_sqnorm(const double* J, double* out)
{
	for (int q = 0; q < 4; ++q) out[q] = 0;
	for (int i = 0; i < NPIX; ++i)
		for (int q = 0; q < 4; ++q) out[q] += J[i * 4 + q] * J[i * 4 + q];
}

/* TrustRegionMinimizer::EvaluateGradientAndJacobian, ceres/internal/ceres/
 * trust_region_minimizer.cc:231-288. */
static int eval_grad_jac(const oproblem* P, const double* x, int iter0,
						 double* x_cost, double* r, double* J, double* scale,
						 double* grad_max_norm)
{
	double g[4] = {0, 0, 0, 0};
	if (!evaluate(P, x, x_cost, r, J)) return 0;
	for (int i = 0; i < NPIX; ++i)
		for (int q = 0; q < 4; ++q) g[q] += J[i * 4 + q] * r[i];
	if (iter0)
	{
		col_sqnorm(J, scale);
		for (int q = 0; q < 4; ++q) scale[q] = 1.0 / (1.0 + sqrt(scale[q]));
	}
	for (int i = 0; i < NPIX; ++i)
		for (int q = 0; q < 4; ++q) J[i * 4 + q] *= scale[q];
	double ng[4], xp[5];
	for (int q = 0; q < 4; ++q) ng[q] = -g[q];
	plus5(x, ng, xp);
	double mx = 0;
	for (int k = 0; k < 5; ++k)
	{
		const double d = fabs(x[k] - xp[k]);
		if (d > mx) mx = d;
	}
	*grad_max_norm = mx;
	return 1;
}

static double norm5(const double* a)
{
	double s = 0;
	for (int k = 0; k < 5; ++k) s += a[k] * a[k];
	return sqrt(s);
}

/* TrustRegionMinimizer::Minimize, trust_region_minimizer.cc:67-121 and the
 * helpers it calls; LevenbergMarquardtStrategy, levenberg_marquardt_strategy.
 * cc:67-166; TrustRegionStepEvaluator, trust_region_step_evaluator.cc:40-114;
 * final cost, solver_utils.h:48-56.  params (5) is updated in place to the
 * lowest-cost accepted iterate, or left alone on FAILURE (solver.cc:428-432). */
static void lm_solve(const oproblem* P, double* params, int max_iter, osummary* S)
{
	static double r[NPIX], J[NPIX * 4];
	double x[5], cand[5], best[5];
	double scale[4] = {1, 1, 1, 1}, diag[4], lmd[4], step[4], delta[4];
	double x_cost = DBL_MAX, minimum_cost = DBL_MAX, x_norm = -1;
	double radius = 1e4, decrease_factor = 2.0;
	int reuse_diagonal = 0, consecutive_invalid = 0;
	double grad_max_norm = 0;
	double final_cost;
	memcpy(x, params, sizeof(x));
	memcpy(best, params, sizeof(best));
	memset(S, 0, sizeof(*S));
	S->termination = 0;

	/* IterationZero */
	S->n_jac_evals++;
	if (!eval_grad_jac(P, x, 1, &x_cost, r, J, scale, &grad_max_norm))
	{
		S->termination = 2;
		S->final_cost = -1.0; /* Summary::initial_cost default */
		return;
	}
	final_cost = x_cost;
	/* step evaluator */
	double se_min = x_cost, se_cur = x_cost, se_ref = x_cost, se_cand = x_cost;
	double se_acc_ref = 0, se_acc_cand = 0;
	int se_nonmono = 0;
	const int se_max_nonmono = 5;

	int iteration = 0;
	int step_successful = 1;
	double iter_cost = x_cost;
	double iter_gmn = grad_max_norm;
	for (;;)
	{
		/* FinalizeIterationAndCheckIfMinimizerCanContinue */
		if (step_successful)
		{
			if (iteration > 0) S->n_successful++;
			if (x_cost < minimum_cost)
			{
				minimum_cost = x_cost;
				memcpy(best, x, sizeof(best));
			}
		}
		else
		{
			S->n_unsuccessful++;
		}
		if (iter_cost < final_cost) final_cost = iter_cost;
		if (iteration >= max_iter) break;                          /* NO_CONVERGENCE */
		if (step_successful && iter_gmn <= 1e-10) { S->termination = 1; break; }
		if (radius <= 1e-32) { S->termination = 1; break; }

		++iteration;
		step_successful = 0;
		iter_gmn = 0;
		iter_cost = 0;

		/* ComputeTrustRegionStep / LevenbergMarquardtStrategy::ComputeStep */
		if (!reuse_diagonal)
		{
			col_sqnorm(J, diag);
			for (int q = 0; q < 4; ++q)
				diag[q] = fmin(fmax(diag[q], 1e-6), 1e32);
		}
		for (int q = 0; q < 4; ++q) lmd[q] = sqrt(diag[q] / radius);
		qr_solve4(J, lmd, r, step);
		reuse_diagonal = 1;
		int valid = all_finite(step, 4);
		double model_cost_change = 0;
		if (valid)
		{
			for (int q = 0; q < 4; ++q) step[q] = -step[q];
			for (int i = 0; i < NPIX; ++i)
			{
				double mr = 0;
				for (int q = 0; q < 4; ++q) mr += J[i * 4 + q] * step[q];
				model_cost_change += mr * (r[i] + mr / 2.0);
			}
			model_cost_change = -model_cost_change;
			valid = model_cost_change > 0.0;
		}
		if (!valid)
		{
			/* HandleInvalidStep */
			if (++consecutive_invalid >= 5)
			{
				S->termination = 2;
				goto done_failure;
			}
			radius /= decrease_factor;
			decrease_factor *= 2.0;
			reuse_diagonal = 1;
			iter_cost = x_cost;
			continue;
		}
		consecutive_invalid = 0;
		for (int q = 0; q < 4; ++q) delta[q] = step[q] * scale[q];

		/* ComputeCandidatePointAndEvaluateCost */
		double cand_cost;
		plus5(x, delta, cand);
		S->n_cost_evals++;
		if (!evaluate(P, cand, &cand_cost, NULL, NULL)) cand_cost = DBL_MAX;

		/* ParameterToleranceReached */
		{
			double d5[5];
			for (int k = 0; k < 5; ++k) d5[k] = x[k] - cand[k];
			if (norm5(d5) <= 1e-8 * (x_norm + 1e-8)) { S->termination = 1; goto done; }
		}
		/* FunctionToleranceReached */
		if (fabs(x_cost - cand_cost) <= 1e-6 * x_cost) { S->termination = 1; goto done; }

		/* IsStepSuccessful / StepQuality */
		double rho;
		if (cand_cost >= DBL_MAX)
			rho = -DBL_MAX;
		else
		{
			const double rd = (se_cur - cand_cost) / model_cost_change;
			const double hd = (se_ref - cand_cost) / (se_acc_ref + model_cost_change);
			rho = rd > hd ? rd : hd;
		}
		if (rho > 1e-3)
		{
			/* HandleSuccessfulStep */
			memcpy(x, cand, sizeof(x));
			x_norm = norm5(x);
			S->n_jac_evals++;
			if (!eval_grad_jac(P, x, 0, &x_cost, r, J, scale, &grad_max_norm))
			{
				S->termination = 2;
				goto done_failure;
			}
			iter_cost = x_cost;
			iter_gmn = grad_max_norm;
			step_successful = 1;
			/* StepAccepted (strategy) */
			{
				const double t = 2.0 * rho - 1.0;
				radius = radius / fmax(1.0 / 3.0, 1.0 - t * t * t);
				radius = fmin(1e16, radius);
				decrease_factor = 2.0;
				reuse_diagonal = 0;
			}
			/* StepAccepted (step evaluator) */
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
			if (se_nonmono == se_max_nonmono)
			{
				se_ref = se_cand;
				se_acc_ref = se_acc_cand;
			}
		}
		else
		{
			/* HandleUnsuccessfulStep */
			radius /= decrease_factor;
			decrease_factor *= 2.0;
			reuse_diagonal = 1;
			iter_cost = cand_cost;
		}
	}
done:
	memcpy(params, best, sizeof(best));
	S->final_cost = final_cost;
	return;
done_failure:
	/* FAILURE: solution not usable, user state restored (solver.cc:428-432) */
	S->final_cost = final_cost;
}

/* ------------------------------------------------------------------------ */
/* tracker::Patch (ft/include/feature_tracker/patch.h:98-130)                */
/* ------------------------------------------------------------------------ */
#define RING 512
typedef struct
{
	int init, lost, track_id;
	double rx, ry; /* patch_.tl() */
	double ix, iy; /* initPoint_ */
	int64_t cur_ts, time_last_update, time_without_update, init_time;
	ebt_event ring[RING]; /* events_: element k-th newest = ring[(tail+size-1-k)&511] */
	int tail, size;
	int num_of_events, counter;
	double nabla[NPIX];
	double mc_nabla[NPIX];   /* motionCompensatedIntegratedNabla_ */
	double pred_nabla[NPIX]; /* predictedNabla_ */
	double cost_map[121];    /* costMap_ (11x11 once Optimizer::drawCostMap ran, zeros before) */
	double flow;    /* flowDir_ (double); read through float getFlow() */
	double warp[4]; /* cos sin tx ty */
	double* traj;   /* (x, y, ts) triples */
	int n_traj, cap_traj;
	double* costs;
	int n_costs, cap_costs;
	int frame_slot;
} opatch;

static void corner_of(const opatch* p, double* x, double* y)
{
	/* Patch::toCorner, ft/src/patch.cpp:166-170 */
	*x = p->rx + (SIDE - 1) / 2.;
	*y = p->ry + (SIDE - 1) / 2.;
}

static void add_traj(opatch* p)
{
	/* Patch::addTrajectoryPosition, patch.cpp:255-258 */
	if (p->n_traj == p->cap_traj)
	{
		p->cap_traj = p->cap_traj ? 2 * p->cap_traj : 64;
		p->traj = (double*)realloc(p->traj, 3 * sizeof(double) * p->cap_traj);
	}
	corner_of(p, &p->traj[3 * p->n_traj], &p->traj[3 * p->n_traj + 1]);
	p->traj[3 * p->n_traj + 2] = (double)p->cur_ts;
	p->n_traj++;
}

/* Patch::Patch + init, patch.cpp:8-35 */
static void patch_init(opatch* p, double cx, double cy, int64_t ts)
{
	memset(p, 0, sizeof(*p));
	p->cur_ts = ts;
	p->rx = cx - EBT_PATCH_EXTENT;
	p->ry = cy - EBT_PATCH_EXTENT;
	p->track_id = -1;
	p->num_of_events = 75;
	corner_of(p, &p->ix, &p->iy);
	p->time_without_update = (int64_t)1e7;
	p->init_time = ts;
	p->warp[0] = 1;
	add_traj(p);
}

static void patch_free(opatch* p)
{
	free(p->traj);
	free(p->costs);
}

static void patch_clone(opatch* dst, const opatch* src)
{
	*dst = *src;
	dst->traj = (double*)malloc(3 * sizeof(double) * (src->cap_traj ? src->cap_traj : 1));
	memcpy(dst->traj, src->traj, 3 * sizeof(double) * src->n_traj);
	dst->costs = (double*)malloc(sizeof(double) * (src->cap_costs ? src->cap_costs : 1));
	memcpy(dst->costs, src->costs, sizeof(double) * src->n_costs);
}

/* cv::Rect2d::contains(Point2i) */
static int in_rect(const opatch* p, int x, int y)
{
	return p->rx <= x && x < p->rx + SIDE && p->ry <= y && y < p->ry + SIDE;
}

/* Patch::addEvent, patch.cpp:37-47 */
static void patch_add_event(opatch* p, const ebt_event* e)
{
	p->ring[(p->tail + p->size) & (RING - 1)] = *e;
	p->size++;
	while (p->size > p->num_of_events)
	{
		p->tail = (p->tail + 1) & (RING - 1);
		p->size--;
	}
	p->counter++;
}

/* Patch::setNumOfEvents, patch.cpp:208-212 */
static void patch_set_num_of_events(opatch* p, uint64_t n)
{
	if (n < 100) n = 100;
	if (n > 300) n = 300;
	p->num_of_events = (int)n;
}

/* Patch::isReady, patch.cpp:296-299 */
static int patch_ready(const opatch* p)
{
	return p->counter >= 30 && p->size >= p->num_of_events;
}

static int32_t to_i32(double v)
{
	/* static_cast<int32_t>(double): truncation; out-of-range is UB in C++ and
	 * yields INT32_MIN on x86-64 (cvttsd2si) */
	if (!(v > -2147483649.0 && v < 2147483648.0)) return INT32_MIN;
	return (int32_t)v;
}

/* Patch::integrateEvents, patch.cpp:65-84 */
static void patch_integrate(opatch* p)
{
	for (int i = 0; i < NPIX; ++i) p->nabla[i] = 0;
	for (int k = 0; k < p->size; ++k)
	{
		const ebt_event* e = &p->ring[(p->tail + p->size - 1 - k) & (RING - 1)];
		if (in_rect(p, e->x, e->y))
		{
			/* frameToPatchCoords: Point2i(double,double) truncates */
			const int px = (int)(e->x - p->rx), py = (int)(e->y - p->ry);
			p->nabla[py * SIDE + px] += (double)e->polarity;
		}
	}
	const ebt_event* front = &p->ring[(p->tail + p->size - 1) & (RING - 1)];
	const ebt_event* back = &p->ring[p->tail];
	p->cur_ts = to_i32((double)(front->t_us + back->t_us) * 0.5);
	p->time_last_update = to_i32((double)back->t_us);
}

/* Patch::updatePatchRect, patch.cpp:49-63; SE2::inverse sophus/se2.hpp:145-148,
 * SO2::inverse = conjugate. */
static void patch_update_rect(opatch* p)
{
	const double c = p->warp[0], s = p->warp[1], tx = p->warp[2], ty = p->warp[3];
	/* invR = (c, -s); t' = invR * (t * -1) */
	const double mx = tx * -1.0, my = ty * -1.0;
	const double itx = c * mx - (-s) * my;
	const double ity = (-s) * mx + c * my;
	/* matrix2x3 of inverse: [c, s; -s, c | itx, ity] */
	const double ncx = c * p->ix + s * p->iy + itx;
	const double ncy = -s * p->ix + c * p->iy + ity;
	p->rx = ncx - EBT_PATCH_EXTENT;
	p->ry = ncy - EBT_PATCH_EXTENT;
}

static int cv_round(double v) { return (int)nearbyint(v); }

/* Patch::integrateMotionCompensatedEvents, patch.cpp:87-130: the window's
 * events moved along the last trajectory segment to the window's mid time.
 * static_cast<Point2i>(Point2d) rounds half-to-even, frameToPatchCoords
 * truncates, containment is tested against the (already updated) rect. */
static void patch_integrate_motion_compensated(opatch* p)
{
	if (p->n_traj < 2 || p->size == 0) return;
	const double* last = &p->traj[3 * (p->n_traj - 1)];
	const double* pre = &p->traj[3 * (p->n_traj - 2)];
	const int64_t last_ts = (int64_t)last[2], pre_ts = (int64_t)pre[2];
	const int64_t mid = p->cur_ts;
	const int64_t half = to_i32((double)(last_ts - pre_ts) * 0.5);
	if (!(last_ts + half >= mid && pre_ts < mid)) return;
	for (int i = 0; i < NPIX; ++i) p->mc_nabla[i] = 0;
	const double dirx = last[0] - pre[0], diry = last[1] - pre[1];
	const double t_dif = (double)(last_ts - pre_ts);
	const double t = (double)mid;
	for (int k = 0; k < p->size; ++k)
	{
		const ebt_event* e = &p->ring[(p->tail + p->size - 1 - k) & (RING - 1)];
		const double sc = (t - (double)e->t_us) / t_dif;
		const double cxd = (double)e->x + dirx * sc, cyd = (double)e->y + diry * sc;
		if (!(fabs(cxd) < 2e9 && fabs(cyd) < 2e9)) continue; /* saturate_cast: never inside */
		const int cx = cv_round(cxd), cy = cv_round(cyd);
		if (in_rect(p, cx, cy))
		{
			const int px = (int)(cx - p->rx), py = (int)(cy - p->ry);
			p->mc_nabla[py * SIDE + px] += (double)e->polarity;
		}
	}
}

/* Patch::warpImage, patch.cpp:132-154: the patch's own template gradients
 * (Patch::gradX_/gradY_, set at creation) warped with cv::warpAffine(flags =
 * WARP_INVERSE_MAP) = nearest neighbour, 10-bit fixed point, cropped to
 * Rect(patch_) (each field rounded half-to-even);
 * predictedNabla = -gx cos(flowDir_) - gy sin(flowDir_).  Unchanged when the
 * rect is not strictly inside the image. */
static void patch_warp_image(opatch* p, const oframe* f)
{
	if (p->rx < 0 || p->ry < 0 || p->rx + SIDE >= f->w || p->ry + SIDE >= f->h) return;
	const double c = p->warp[0], s = p->warp[1], tx = p->warp[2], ty = p->warp[3];
	const int x0 = cv_round(p->rx), y0 = cv_round(p->ry);
	const double a = -cos(p->flow), b = -sin(p->flow);
	for (int yy = 0; yy < SIDE; ++yy)
		for (int xx = 0; xx < SIDE; ++xx)
		{
			const int x = x0 + xx, y = y0 + yy;
			double vx = 0, vy = 0;
			if (x >= 0 && x < f->w && y >= 0 && y < f->h) /* crop of the warped full frame */
			{
				const int X0 = cv_round(c * x * 1024) + cv_round((-s * y + tx) * 1024) + 512;
				const int Y0 = cv_round(s * x * 1024) + cv_round((c * y + ty) * 1024) + 512;
				const int xi = X0 >> 10, yi = Y0 >> 10;
				if (xi >= 0 && xi < f->w && yi >= 0 && yi < f->h)
				{
					vx = f->g[2 * (yi * f->w + xi)];
					vy = f->g[2 * (yi * f->w + xi) + 1];
				}
			}
			p->pred_nabla[yy * SIDE + xx] = vx * a + vy * b;
		}
}

/* OptimizerParams::drawCostMap (ft/include/feature_tracker/optimizer.h:18) for
 * every optimisation that follows. */
static int g_draw_cost_map = 0;
void oracle_set_draw_cost_map(int on) { g_draw_cost_map = on; }

/* Optimizer::drawCostMap, ft/src/optimizer.cpp:32-60: the functor of the solve
 * (old rect, normalised patch) evaluated on an 11x11 grid of integer translation
 * offsets around the SOLVED warp; the pose is rebuilt from its angle
 * (Pose2d(pose.log().z(), t): SO2::log = atan2, SO2(theta) = (cos, sin)); the
 * flow is the float getFlow(); the map holds cv::norm(residuals, NORM_L2) of the
 * uncorrected residuals r_i = u_i / sqrt(N) + m_i (optimizer_cost.h:29-45). */
static void draw_cost_map(opatch* p, const oproblem* P)
{
	const double theta = atan2(p->warp[1], p->warp[0]);
	const double cs = cos(theta), sn = sin(theta);
	const double phi = (double)(float)p->flow;
	const double vx = cos(phi), vy = sin(phi);
	const int W = P->frame->w, H = P->frame->h;
	for (int ox = -5; ox <= 5; ++ox)
		for (int oy = -5; oy <= 5; ++oy)
		{
			const double tx = (double)(float)ox + p->warp[2], ty = (double)(float)oy + p->warp[3];
			double u[NPIX], N = 1e-5;
			for (int y = 0; y < SIDE; ++y)
				for (int xx = 0; xx < SIDE; ++xx)
				{
					const int i = xx + SIDE * y;
					const double px = xx + P->tlx, py = y + P->tly;
					const double wx = cs * px + -sn * py + tx;
					const double wy = sn * px + cs * py + ty;
					u[i] = 0;
					if (wx >= W || wy >= H || wx < 0.0 || wy < 0.0) continue;
					double g[2];
					bicubic(P->frame, wy, wx, g, NULL, NULL);
					u[i] = g[0] * vx + g[1] * vy;
					N += u[i] * u[i];
				}
			const double sq = sqrt(N);
			double s = 0;
			for (int i = 0; i < NPIX; ++i)
			{
				const double r = u[i] / sq + P->m[i];
				s += r * r;
			}
			p->cost_map[(oy + 5) * 11 + (ox + 5)] = sqrt(s);
		}
}

/* Optimizer::optimize, ft/src/optimizer.cpp:62-206. */
static void optimize_patch(opatch* p, const oframe* frame, double huber_a,
						   double threshold, int max_iter, osummary* S)
{
	patch_integrate(p);
	/* Patch::getNormalizedIntegratedNabla, patch.cpp:156-159: OpenCV evaluates
	 * Mat / s as Mat * (1/s) */
	double m[NPIX], nrm = 0;
	for (int i = 0; i < NPIX; ++i) nrm += p->nabla[i] * p->nabla[i];
	nrm = sqrt(nrm);
	const double inv = 1.0 / nrm;
	for (int i = 0; i < NPIX; ++i) m[i] = p->nabla[i] * inv;

	double x[5] = {p->warp[0], p->warp[1], p->warp[2], p->warp[3],
				   (double)(float)p->flow}; /* getFlow() returns float */
	oproblem P = {frame, m, p->rx, p->ry, huber_a};
	lm_solve(&P, x, max_iter, S);

	/* addFinalCost */
	if (p->n_costs == p->cap_costs)
	{
		p->cap_costs = p->cap_costs ? 2 * p->cap_costs : 64;
		p->costs = (double*)realloc(p->costs, sizeof(double) * p->cap_costs);
	}
	p->costs[p->n_costs++] = S->final_cost;

	/* cost filter, optimizer.cpp:136-157: reads costs[n-6+i], i = 0..4.  At
	 * n == 5 the reference reads index -1 (UB); restated as "needs n >= 6". */
	if (p->n_costs >= 6)
	{
		double v[5];
		for (int i = 0; i < 5; ++i) v[i] = p->costs[p->n_costs - 6 + i];
		for (int i = 1; i < 5; ++i)
			for (int j = i; j > 0 && v[j] < v[j - 1]; --j)
			{
				const double t = v[j];
				v[j] = v[j - 1];
				v[j - 1] = t;
			}
		if (v[2] > threshold)
		{
			p->lost = 1;
			return;
		}
	}
	p->flow = fmod(x[4], 2 * M_PI);
	memcpy(p->warp, x, 4 * sizeof(double));
	p->init = 1;
	double ox, oy, nx, ny;
	corner_of(p, &ox, &oy);
	patch_update_rect(p);
	corner_of(p, &nx, &ny);
	const double mv = sqrt((nx - ox) * (nx - ox) + (ny - oy) * (ny - oy));
	p->time_without_update = (int64_t)(1e6 / fmax(1e-1, mv));
	add_traj(p);
	patch_integrate_motion_compensated(p);
	p->counter = 0; /* resetBatch */
	if (g_draw_cost_map) draw_cost_map(p, &P); /* optimizer.cpp:191-197 */
}

/* ------------------------------------------------------------------------ */
/* tracker::FeatureDetector — tracking part (ft/src/feature_detector.cpp)    */
/* ------------------------------------------------------------------------ */
#define MAX_FRAMES 4096 /* optimizers_ is an unbounded map in the reference */
typedef struct
{
	int w, h, max_patches;
	double association_distance;
	int init_num_events;
	double huber_a, threshold;
	int max_iter;
	oframe frames[MAX_FRAMES]; /* optimizers_ keyed by image ts */
	int latest;                /* slot of gradX_/gradY_ (latest frame) */
	opatch** patches;          /* patches_ (list order) */
	int n_patches;
	opatch** archived;
	int n_archived, cap_archived;
	int next_track_id;
	int image_counter;
	ebt_opt_record* records;
	int64_t n_records, cap_records;
	int64_t event_index;
	int draw_images; /* DetectorParams::drawImages */
} otracker;

/* FeatureDetector::updateNumOfEvents, feature_detector.cpp:666-712, with
 * cv::warpAffine(flags = WARP_INVERSE_MAP) = nearest neighbour, 10-bit fixed
 * point, evaluated on the cropped rect only. */
static void update_num_of_events(otracker* T, opatch* p)
{
	const oframe* f = &T->frames[T->latest];
	const double cx = (p->rx + (p->rx + SIDE)) * 0.5;
	const double cy = (p->ry + (p->ry + SIDE)) * 0.5;
	if (cx <= 5 || cy <= 5 || cx >= f->w - 5 || cy >= f->h - 5)
	{
		p->lost = 1;
		return;
	}
	if (p->rx < 0 || p->ry < 0 || p->rx + SIDE >= f->w || p->ry + SIDE >= f->h)
	{
		patch_set_num_of_events(p, (uint64_t)T->init_num_events);
		return;
	}
	const double c = p->warp[0], s = p->warp[1], tx = p->warp[2], ty = p->warp[3];
	const int x0 = cv_round(p->rx), y0 = cv_round(p->ry);
	const float flow = (float)p->flow;
	const double a = 0.6 * (double)cosf(flow);
	const double b = 0.6 * (double)sinf(flow);
	double acc = 0;
	for (int y = y0; y < y0 + SIDE; ++y)
		for (int x = x0; x < x0 + SIDE; ++x)
		{
			const int X0 = cv_round(c * x * 1024) + cv_round((-s * y + tx) * 1024) + 512;
			const int Y0 = cv_round(s * x * 1024) + cv_round((c * y + ty) * 1024) + 512;
			const int xi = X0 >> 10, yi = Y0 >> 10;
			double vx = 0, vy = 0;
			if (xi >= 0 && xi < f->w && yi >= 0 && yi < f->h)
			{
				vx = f->g[2 * (yi * f->w + xi)];
				vy = f->g[2 * (yi * f->w + xi) + 1];
			}
			acc += fabs(vx * a + vy * b);
		}
	patch_set_num_of_events(p, (uint64_t)acc);
}

static int find_frame(const otracker* T, int64_t ts)
{
	for (int i = 0; i < MAX_FRAMES; ++i)
		if (T->frames[i].alive && T->frames[i].ts == ts) return i;
	return -1;
}

void* oracle_create(int width, int height, int max_patches)
{
	otracker* T = (otracker*)calloc(1, sizeof(otracker));
	T->w = width;
	T->h = height;
	T->max_patches = max_patches;
	T->association_distance = 5;
	T->init_num_events = 75;
	T->huber_a = 0.3;
	T->threshold = 0.6;
	T->max_iter = 10;
	T->patches = (opatch**)calloc((size_t)max_patches + 1, sizeof(opatch*));
	return T;
}

void oracle_destroy(void* h)
{
	otracker* T = (otracker*)h;
	for (int i = 0; i < T->n_patches; ++i) { patch_free(T->patches[i]); free(T->patches[i]); }
	for (int i = 0; i < T->n_archived; ++i) { patch_free(T->archived[i]); free(T->archived[i]); }
	for (int i = 0; i < MAX_FRAMES; ++i) free(T->frames[i].g);
	free(T->patches);
	free(T->archived);
	free(T->records);
	free(T);
}

/* extractPatches (corners injected) + associatePatches, feature_detector.cpp:
 * 543-566, 630-664. */
int oracle_new_frame_begin(void* h, const uint8_t* gray, int64_t ts,
						   const double* cornersXY, int n_corners)
{
	otracker* T = (otracker*)h;
	int slot = find_frame(T, ts);
	if (slot < 0)
		for (int i = 0; i < MAX_FRAMES && slot < 0; ++i)
			if (!T->frames[i].alive) slot = i;
	if (slot < 0) return -1;
	oframe* f = &T->frames[slot];
	if (!f->g) f->g = (double*)malloc(sizeof(double) * 2 * (size_t)T->w * T->h);
	f->w = T->w;
	f->h = T->h;
	f->ts = ts;
	f->users = 0;
	f->alive = 1;
	log_sobel(gray, T->w, T->h, f->g);
	T->latest = slot;

	int* new_ids = (int*)malloc(sizeof(int) * (size_t)(n_corners + 1));
	for (int j = 0; j < n_corners; ++j) new_ids[j] = -1;
	for (int i = 0; i < T->n_patches; ++i)
	{
		opatch* p = T->patches[i];
		double cx, cy;
		corner_of(p, &cx, &cy);
		for (int j = 0; j < n_corners; ++j)
		{
			const double dx = cx - cornersXY[2 * j], dy = cy - cornersXY[2 * j + 1];
			if (new_ids[j] == -1 && sqrt(dx * dx + dy * dy) < T->association_distance)
			{
				new_ids[j] = p->track_id;
				break;
			}
		}
		p->cur_ts = ts; /* setTs */
		add_traj(p);
	}
	for (int j = 0; j < n_corners; ++j)
	{
		if (new_ids[j] == -1 && T->n_patches < T->max_patches)
		{
			opatch* p = (opatch*)malloc(sizeof(opatch));
			patch_init(p, cornersXY[2 * j], cornersXY[2 * j + 1], ts);
			p->track_id = T->next_track_id++;
			p->frame_slot = slot;
			T->patches[T->n_patches++] = p;
			f->users++;
		}
	}
	free(new_ids);
	if (T->image_counter < 2) T->image_counter++;
	return slot;
}

/* FlowEstimator::getFlowPatches query side, ft/src/flow_estimator.cpp:27-39 */
int oracle_get_uninit(void* h, int32_t* ids, double* cornersXY, int cap)
{
	otracker* T = (otracker*)h;
	int n = 0;
	if (T->image_counter != 2) return 0;
	for (int i = 0; i < T->n_patches; ++i)
	{
		opatch* p = T->patches[i];
		if (p->init) continue;
		if (n < cap)
		{
			ids[n] = p->track_id;
			corner_of(p, &cornersXY[2 * n], &cornersXY[2 * n + 1]);
		}
		++n;
	}
	return n;
}

/* loop body of getFlowPatches, flow_estimator.cpp:41-79, LK result injected */
void oracle_set_flow(void* h, int32_t track_id, int found, float next_x, float next_y)
{
	otracker* T = (otracker*)h;
	for (int i = 0; i < T->n_patches; ++i)
	{
		opatch* p = T->patches[i];
		if (p->track_id != track_id || p->init) continue;
		if (!found)
		{
			p->lost = 1;
			return;
		}
		double cx, cy;
		corner_of(p, &cx, &cy);
		const double dx = next_x - cx, dy = next_y - cy;
		p->warp[0] = 1;
		p->warp[1] = 0;
		p->warp[2] = -dx;
		p->warp[3] = -dy;
		p->flow = atan2(dy, dx);
		p->init = 1;
		p->time_without_update = (int64_t)(1e6 / fmax(1e-1, sqrt(dx * dx + dy * dy)));
		if (cx <= 5 || cy <= 5 || cx >= T->w - 5 || cy >= T->h - 5) p->lost = 1;
		return;
	}
}

/* newImage tail, feature_detector.cpp:499-535 */
void oracle_new_frame_end(void* h)
{
	otracker* T = (otracker*)h;
	for (int i = 0; i < T->n_patches; ++i) update_num_of_events(T, T->patches[i]);
	if (T->draw_images) /* feature_detector.cpp:504-510 */
		for (int i = 0; i < T->n_patches; ++i)
			patch_warp_image(T->patches[i], &T->frames[T->patches[i]->frame_slot]);
	int k = 0;
	for (int i = 0; i < T->n_patches; ++i)
	{
		opatch* p = T->patches[i];
		if (p->lost)
		{
			if (T->n_archived == T->cap_archived)
			{
				T->cap_archived = T->cap_archived ? 2 * T->cap_archived : 64;
				T->archived = (opatch**)realloc(T->archived, sizeof(opatch*) * T->cap_archived);
			}
			T->archived[T->n_archived++] = p;
			T->frames[p->frame_slot].users--;
		}
		else
			T->patches[k++] = p;
	}
	T->n_patches = k;
	for (int i = 0; i < MAX_FRAMES; ++i)
		if (T->frames[i].alive && T->frames[i].users == 0) T->frames[i].alive = 0;
}

static void push_record(otracker* T, const opatch* p, const osummary* S)
{
	if (T->n_records == T->cap_records)
	{
		T->cap_records = T->cap_records ? 2 * T->cap_records : 1024;
		T->records = (ebt_opt_record*)realloc(T->records, sizeof(ebt_opt_record) * T->cap_records);
	}
	ebt_opt_record* r = &T->records[T->n_records++];
	memset(r, 0, sizeof(*r));
	r->track_id = p->track_id;
	r->n_successful = S->n_successful;
	r->n_unsuccessful = S->n_unsuccessful;
	r->termination = S->termination;
	r->lost = p->lost;
	r->num_of_events = p->num_of_events;
	r->ts = p->cur_ts;
	r->event_index = T->event_index;
	corner_of(p, &r->x, &r->y);
	r->final_cost = S->final_cost;
	memcpy(r->warp, p->warp, sizeof(r->warp));
	r->flow_dir = (double)(float)p->flow;
}

/* FeatureDetector::updatePatches, feature_detector.cpp:585-619, per event */
int64_t oracle_push_events(void* h, const ebt_event* ev, int64_t n)
{
	otracker* T = (otracker*)h;
	int64_t n_opt = 0;
	for (int64_t i = 0; i < n; ++i, ++T->event_index)
	{
		for (int k = 0; k < T->n_patches; ++k)
		{
			opatch* p = T->patches[k];
			if (p->lost) continue;
			if (in_rect(p, ev[i].x, ev[i].y)) patch_add_event(p, &ev[i]);
			if (patch_ready(p) && p->init)
			{
				osummary S;
				optimize_patch(p, &T->frames[p->frame_slot], T->huber_a, T->threshold,
							   T->max_iter, &S);
				update_num_of_events(T, p);
				patch_warp_image(p, &T->frames[p->frame_slot]); /* feature_detector.cpp:615 */
				push_record(T, p, &S);
				++n_opt;
			}
		}
	}
	return n_opt;
}

static void fill_track(const opatch* p, ebt_track* t)
{
	memset(t, 0, sizeof(*t));
	t->track_id = p->track_id;
	t->lost = p->lost;
	t->init = p->init;
	t->num_of_events = p->num_of_events;
	t->counter = p->counter;
	t->window_size = p->size;
	t->n_optimizations = p->n_costs;
	t->n_trajectory = p->n_traj;
	t->rect_x = p->rx;
	t->rect_y = p->ry;
	t->init_x = p->ix;
	t->init_y = p->iy;
	memcpy(t->warp, p->warp, sizeof(t->warp));
	t->flow_dir = p->init ? (double)(float)p->flow : 0;
	t->last_cost = p->n_costs ? p->costs[p->n_costs - 1] : 0;
	t->current_ts = p->cur_ts;
	t->time_last_update = p->time_last_update;
	t->time_without_update = p->time_without_update;
	t->init_time = p->init_time;
}

int oracle_num_patches(void* h) { return ((otracker*)h)->n_patches; }
int oracle_num_archived(void* h) { return ((otracker*)h)->n_archived; }
int oracle_get_tracks(void* h, ebt_track* out, int cap, int archived)
{
	otracker* T = (otracker*)h;
	opatch** list = archived ? T->archived : T->patches;
	const int n = archived ? T->n_archived : T->n_patches;
	for (int i = 0; i < n && i < cap; ++i) fill_track(list[i], &out[i]);
	return n;
}
int64_t oracle_num_records(void* h) { return ((otracker*)h)->n_records; }
int64_t oracle_get_records(void* h, ebt_opt_record* out, int64_t cap)
{
	otracker* T = (otracker*)h;
	const int64_t n = cap < T->n_records ? cap : T->n_records;
	memcpy(out, T->records, (size_t)n * sizeof(ebt_opt_record));
	return n;
}
void oracle_clear_records(void* h) { ((otracker*)h)->n_records = 0; }

static opatch* find_patch(otracker* T, int32_t id)
{
	for (int i = 0; i < T->n_patches; ++i)
		if (T->patches[i]->track_id == id) return T->patches[i];
	for (int i = 0; i < T->n_archived; ++i)
		if (T->archived[i]->track_id == id) return T->archived[i];
	return NULL;
}
int oracle_get_patch(void* h, int32_t id, int which, double* out625)
{
	opatch* p = find_patch((otracker*)h, id);
	/* which: 0 integratedNabla_, 1 motionCompensatedIntegratedNabla_, 2 predictedNabla_, 3 costMap_ */
	if (!p || which < 0 || which > 3) return -1;
	if (which == 3)
	{
		memcpy(out625, p->cost_map, sizeof(p->cost_map));
		return 121;
	}
	memcpy(out625, which == 0 ? p->nabla : which == 1 ? p->mc_nabla : p->pred_nabla, sizeof(p->nabla));
	return NPIX;
}
void oracle_set_draw_images(void* h, int on) { ((otracker*)h)->draw_images = on; }
int oracle_get_trajectory(void* h, int32_t id, double* out, int cap)
{
	opatch* p = find_patch((otracker*)h, id);
	if (!p) return -1;
	for (int i = 0; i < p->n_traj && i < cap; ++i) memcpy(&out[3 * i], &p->traj[3 * i], 3 * sizeof(double));
	return p->n_traj;
}
int oracle_get_gradients(void* h, double* gx, double* gy)
{
	otracker* T = (otracker*)h;
	const oframe* f = &T->frames[T->latest];
	for (int i = 0; i < f->w * f->h; ++i)
	{
		gx[i] = f->g[2 * i];
		gy[i] = f->g[2 * i + 1];
	}
	return f->w * f->h;
}

/* ------------------------------------------------------------------------ */
/* Batched entry: F independent feature-patch problems                       */
/* ------------------------------------------------------------------------ */
typedef struct
{
	int w, h, n;
	oframe f[MAX_FRAMES];
} oframes;

void* oracle_frames_create(int width, int height)
{
	oframes* F = (oframes*)calloc(1, sizeof(oframes));
	F->w = width;
	F->h = height;
	return F;
}
void oracle_frames_destroy(void* f)
{
	oframes* F = (oframes*)f;
	for (int i = 0; i < F->n; ++i) free(F->f[i].g);
	free(F);
}
int oracle_frames_add(void* f, const uint8_t* gray)
{
	oframes* F = (oframes*)f;
	if (F->n >= MAX_FRAMES) return -1;
	oframe* fr = &F->f[F->n];
	fr->w = F->w;
	fr->h = F->h;
	fr->alive = 1;
	fr->g = (double*)malloc(sizeof(double) * 2 * (size_t)F->w * F->h);
	log_sobel(gray, F->w, F->h, fr->g);
	return F->n++;
}
int oracle_frames_get(void* f, int slot, double* gx, double* gy)
{
	oframes* F = (oframes*)f;
	for (int i = 0; i < F->w * F->h; ++i)
	{
		gx[i] = F->f[slot].g[2 * i];
		gy[i] = F->f[slot].g[2 * i + 1];
	}
	return 0;
}

static void problem_to_patch(opatch* p, const ebt_problem* pr, const ebt_event* events)
{
	patch_init(p, pr->init_x, pr->init_y, 0);
	memcpy(p->warp, pr->warp, sizeof(p->warp));
	p->flow = pr->flow_dir;
	p->init = 1;
	patch_update_rect(p);
	patch_set_num_of_events(p, (uint64_t)pr->n_events);
	for (int i = pr->n_events - 1; i >= 0; --i) patch_add_event(p, &events[pr->event_offset + i]);
}

/* nabla (optional): n x 625 doubles.  iters (optional): n x 2 ints
 * (jacobian evaluations, cost-only evaluations). */
static double* g_cost_maps_out = NULL;
double oracle_optimize_batch(void* f, const ebt_problem* problems, int64_t n, const ebt_event* events,
							 ebt_solution* out, double* nabla, int32_t* evals);
/* oracle_optimize_batch with drawCostMap on: cost_maps = n x 121 doubles. */
double oracle_optimize_batch_cm(void* f, const ebt_problem* problems, int64_t n,
								const ebt_event* events, ebt_solution* out,
								double* nabla, double* cost_maps)
{
	const int was = g_draw_cost_map;
	g_draw_cost_map = 1;
	g_cost_maps_out = cost_maps;
	oracle_optimize_batch(f, problems, n, events, out, nabla, NULL);
	g_cost_maps_out = NULL;
	g_draw_cost_map = was;
	return 0;
}

double oracle_optimize_batch(void* f, const ebt_problem* problems, int64_t n,
							 const ebt_event* events, ebt_solution* out,
							 double* nabla, int32_t* evals)
{
	oframes* F = (oframes*)f;
	opatch* p = (opatch*)malloc(sizeof(opatch));
	for (int64_t i = 0; i < n; ++i)
	{
		osummary S;
		problem_to_patch(p, &problems[i], events);
		optimize_patch(p, &F->f[problems[i].frame_slot], 0.3, 0.6, 10, &S);
		ebt_solution* s = &out[i];
		memset(s, 0, sizeof(*s));
		memcpy(s->warp, p->warp, sizeof(s->warp));
		s->flow_dir = (double)(float)p->flow;
		s->final_cost = S.final_cost;
		corner_of(p, &s->x, &s->y);
		s->n_successful = S.n_successful;
		s->n_unsuccessful = S.n_unsuccessful;
		s->termination = S.termination;
		s->ts = p->cur_ts;
		int in = 0;
		for (int k = 0; k < NPIX; ++k)
		{
			if (nabla) nabla[i * NPIX + k] = p->nabla[k];
			in += (int)fabs(p->nabla[k]);
		}
		s->n_passes = -1;
		(void)in;
		if (g_cost_maps_out) memcpy(g_cost_maps_out + i * 121, p->cost_map, sizeof(p->cost_map));
		if (evals)
		{
			evals[2 * i] = S.n_jac_evals;
			evals[2 * i + 1] = S.n_cost_evals;
		}
		patch_free(p);
	}
	free(p);
	return 0;
}

/* residuals + local Jacobian (625x4) of the UNcorrected functor (Huber off:
 * huber_a = +inf) for the analytic-Jacobian check against Ceres autodiff. */
int oracle_eval_functor(void* f, int slot, const double* m625, double rect_x,
						double rect_y, const double* warp4, double flow,
						double* residuals, double* j_local4)
{
	oframes* F = (oframes*)f;
	oproblem P = {&F->f[slot], m625, rect_x, rect_y, 1e300};
	double x[5] = {warp4[0], warp4[1], warp4[2], warp4[3], flow};
	double cost;
	return evaluate(&P, x, &cost, residuals, j_local4) ? 0 : -1;
}

/* Patch.integrateEventsTest body (ft/test/patch_test.cpp:35-60), extent 12 */
int oracle_patch_integrate(double cx, double cy, const ebt_event* ev, int n,
						   double* out, int64_t* front_back_ts)
{
	opatch* p = (opatch*)malloc(sizeof(opatch));
	patch_init(p, cx, cy, 0);
	for (int i = 0; i < n; ++i) patch_add_event(p, &ev[i]);
	patch_integrate(p);
	memcpy(out, p->nabla, sizeof(p->nabla));
	if (front_back_ts)
	{
		front_back_ts[0] = p->ring[(p->tail + p->size - 1) & (RING - 1)].t_us;
		front_back_ts[1] = p->ring[p->tail].t_us;
		front_back_ts[2] = p->cur_ts;
		front_back_ts[3] = patch_ready(p);
	}
	patch_free(p);
	free(p);
	return NPIX;
}

/* cv::warpAffine nearest-neighbour index map (the rule update_num_of_events
 * uses) for a 2x3 dst->src matrix; pinned against cv2 in tests/golden. */
void oracle_warp_index_map(const double* M6, int w, int h, int32_t* xi, int32_t* yi)
{
	for (int y = 0; y < h; ++y)
		for (int x = 0; x < w; ++x)
		{
			const int X0 = cv_round(M6[0] * x * 1024) + cv_round((M6[1] * y + M6[2]) * 1024) + 512;
			const int Y0 = cv_round(M6[3] * x * 1024) + cv_round((M6[4] * y + M6[5]) * 1024) + 512;
			xi[y * w + x] = X0 >> 10;
			yi[y * w + x] = Y0 >> 10;
		}
}

/* std::cos(float) / std::sin(float) as the host libm evaluates them
 * (feature_detector.cpp:702-704), for the GPU float-rounding check. */
void oracle_cosf_sinf(const float* in, int n, float* c, float* s)
{
	for (int i = 0; i < n; ++i)
	{
		c[i] = cosf(in[i]);
		s[i] = sinf(in[i]);
	}
}

/* ------------------------------------------------------------------------ */
/* Tracker -> VO hand-off: common::CameraModel<double>::unproject            */
/* (common/include/common/camera_model.h:79-114; getRadialDistortion :42-47, */
/* getTangentialDistortion :35-40), built with -ffp-contract=off.            */
/* ------------------------------------------------------------------------ */
static double cam_tangential(double p1, double p2, double x, double y, double r2)
{
	return 2.0 * p1 * x * y + p2 * (r2 + 2.0 * x * x);
}
void oracle_unproject(const ebt_camera* cam, const double* xy, int n, double* out3)
{
	for (int i = 0; i < n; ++i)
	{
		const double xd = (xy[2 * i] - cam->cx) / cam->fx, yd = (xy[2 * i + 1] - cam->cy) / cam->fy;
		double xo = xd, yo = yd;
		for (int it = 0; it < 10; ++it)
		{
			const double r2 = xo * xo + yo * yo;
			const double radial = 1.0 + cam->k1 * r2 + cam->k2 * r2 * r2;
			const double dx = cam_tangential(cam->p1, cam->p2, xo, yo, r2);
			const double dy = cam_tangential(cam->p2, cam->p1, yo, xo, r2);
			xo = (xd - dx) / radial;
			yo = (yd - dy) / radial;
		}
		const double norm = sqrt(xo * xo + yo * yo + 1.0);
		out3[3 * i] = xo / norm;
		out3[3 * i + 1] = yo / norm;
		out3[3 * i + 2] = 1.0 / norm;
	}
}
/* CameraModel::project (camera_model.h:49-77), for the reference's round-trip test */
void oracle_project(const ebt_camera* cam, const double* xyz, int n, double* out2)
{
	for (int i = 0; i < n; ++i)
	{
		const double xp = xyz[3 * i] / xyz[3 * i + 2], yp = xyz[3 * i + 1] / xyz[3 * i + 2];
		const double r2 = xp * xp + yp * yp;
		const double radial = 1.0 + cam->k1 * r2 + cam->k2 * r2 * r2;
		const double xD = xp * radial + cam_tangential(cam->p1, cam->p2, xp, yp, r2);
		const double yD = yp * radial + cam_tangential(cam->p2, cam->p1, yp, xp, r2);
		out2[2 * i] = cam->fx * xD + cam->cx;
		out2[2 * i + 1] = cam->fy * yD + cam->cy;
	}
}
