// ebt_api.cu — host side of the C-ABI declared in include/ebt.h.
//
// Owns device memory, the per-frame host logic of tracker::FeatureDetector
// (association, archiving, flow injection; /root/reference/implementation/
// feature_tracker/src/feature_detector.cpp:493-566, 630-664) and launches the
// kernels of ebt_kernels.cu.  No CPU implementation of the hot path lives
// here: every entry point that computes does so on the GPU or fails.
#include <cuda_runtime.h>

#include <algorithm>
#include <cmath>
#include <cstddef>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <ctime>
#include <string>
#include <unordered_map>
#include <vector>

#include "../../include/ebt.h"
#include "ebt_device_types.h"
#include "ebt_frontend.h"
#include "ebt_ingest.h"
#include "ebt_kernels.h"

namespace
{
std::string g_create_error;

struct HostTrack
{
	std::vector<double> traj;   // (x, y, ts) triples: Patch::trajectory_
	std::vector<double> costs;  // Patch::finalCosts_
};

struct Frame
{
	double2* grid = nullptr;
	int64_t ts = 0;
	long long users = 0;  // Optimizer::used_
	bool alive = false;
};

enum
{
	K_TRACK = 0,
	K_BATCH = 1,
	K_GRAD = 2,
	K_NUMEV = 3,
	K_CORNERS = 4,
	K_PYRAMID = 5,
	K_LK = 6,
	K_PARSE = 7,
	K_COUNT = 8
};

// Device state of the per-frame re-arm kernels (ebt_frontend.cu): the image
// pyramids of the previous and the current frame (FlowEstimator::previousImage_
// / currentImage_), Scharr derivatives of the previous one, Harris scratch.
struct Frontend
{
	bool ready = false;
	ebt_frontend_params prm;
	int levels = 0;
	int w[EBT_LK_MAX_LEVELS], h[EBT_LK_MAX_LEVELS];
	unsigned char* img[2][EBT_LK_MAX_LEVELS] = {};
	int built[2] = {0, 0};  // pyramid levels valid beyond level 0
	bool have[2] = {false, false};
	int cur = 0;  // img[cur] = currentImage_, img[cur ^ 1] = previousImage_
	short2* deriv[EBT_LK_MAX_LEVELS] = {};
	unsigned char* scratch = nullptr;  // image of a stand-alone ebt_detect_features call
	float* planes = nullptr;  // 3 product planes
	float* resp = nullptr;
	unsigned* max_key = nullptr;
	int* n_cand = nullptr;
	ebt_corner_candidate* cand = nullptr;
	int cand_cap = 0;
	ebt_corner_candidate* h_cand = nullptr;  // pinned
	int* h_ncand = nullptr;                   // pinned
	float* d_pts = nullptr;                   // pts | next, lk_cap points each
	unsigned char* d_status = nullptr;
	float* h_pts = nullptr;  // pinned, same layout
	double* d_land = nullptr;  // hand-off scratch: n x (2 in | 2 corner + 3 bearing) doubles + n ids
	size_t land_cap = 0;
	unsigned char* h_status = nullptr;
	int lk_cap = 0;
};
}  // namespace

struct ebt_ctx
{
	ebt_params prm;
	DevParams dp;
	int device = 0;
	cudaStream_t stream = nullptr;
	std::string err;

	std::vector<Frame> frames = std::vector<Frame>(EBT_INIT_FRAMES);  // grows on demand (allocFrameSlot)
	int latest = 0;
	const double2** d_frame_ptrs = nullptr;  // device copy of frames[i].grid: FrameTable::grid
	size_t frame_ptrs_cap = 0;
	bool frame_ptrs_dirty = true;
	std::vector<const double2*> h_frame_ptrs;

	std::vector<DevPatch> h;        // FeatureDetector::patches_ (list order)
	std::vector<DevPatch> arch;     // FeatureDetector::archivedPatches_
	// Per-track histories (Patch::trajectory_, finalCosts_), indexed by track
	// id.  They are built lazily: pushes and frames only append to the journal
	// (counts live in DevPatch::n_costs / n_traj); materialise() replays it in
	// order when a history is asked for.
	std::vector<HostTrack> all_tracks;
	struct ImgSample { int32_t id; double x, y, ts; };
	struct Pending { int kind; size_t a, b; };  // 0: img_samples[a,b)  1: records[a,b)
	std::vector<ImgSample> img_samples;
	std::vector<Pending> journal;
	std::vector<std::vector<int>> arch_nabla;
	std::vector<std::vector<int>> arch_mc;       // products on: motionCompensatedIntegratedNabla_
	std::vector<std::vector<double>> arch_pred;  // products on: predictedNabla_
	std::vector<std::vector<double>> arch_cost;  // cost map on: costMap_ (121)
	std::vector<int> free_slots;
	std::vector<int> assoc_start, assoc_items, assoc_fill;  // corner buckets of associatePatches
	bool host_dirty = true;  // h holds changes the device has not seen
	int next_track_id = 0;
	std::vector<double> corners;  // FeatureDetector::corners_ of the latest frame
	int image_counter = 0;  // FlowEstimator::imageCounter_
	int64_t event_index = 0;

	DevPatch* d_patches = nullptr;
	ebt_event* d_rings = nullptr;
	int* d_nablas = nullptr;
	int* d_mc_nablas = nullptr;       // allocated by ebt_set_products
	double* d_pred_nablas = nullptr;
	double* d_cost_maps = nullptr;    // allocated by ebt_set_cost_map: [store_slot][121]
	double* d_bcost = nullptr;        // cost maps of the last batched solve: [problem][121]
	size_t cap_bcost = 0, n_bcost = 0;
	ebt_opt_record* d_records = nullptr;
	unsigned long long* d_nrec = nullptr;
	int* d_queue = nullptr;  // work-queue cursor of the shared-pixel-warp stream kernel
	int stream_kernel = 1;   // ebt_set_stream_kernel
	int sm_count = 0;
	// launch scratch of a (multi-stream) push on the shared-pixel-warp kernel, owned by ctxs[0]
	StreamDesc* d_descs = nullptr;
	StreamDesc* h_descs = nullptr;  // pinned
	size_t descs_cap = 0;
	WorkItem* d_work = nullptr;
	WorkItem* h_work = nullptr;  // pinned
	size_t work_cap = 0;
	std::vector<int> work_load, work_idx;
	std::vector<WorkItem> work_tmp;
	cudaEvent_t join_event = nullptr;  // orders this context's own stream before a multi-stream launch
	int shard_index = 0, shard_count = 1;  // ebt_set_shard: this context tracks track ids = index (mod count)
	double opt_per_event = 0.25;       // optimisations per stream event of the last piece (bounds the next piece)
	// ebt_replay: second CUDA stream, ping-pong device buffers and pinned staging of the pipelined event upload
	cudaStream_t copy_stream = nullptr;
	cudaEvent_t uploaded[2] = {nullptr, nullptr};
	cudaEvent_t chunk_done[2][2] = {{nullptr, nullptr}, {nullptr, nullptr}};
	ebt_event* d_replay[2] = {nullptr, nullptr};
	ebt_event* h_replay[2] = {nullptr, nullptr};
	size_t replay_cap = 0;
	unsigned char* arch_stage = nullptr;  // pinned: products of the patches archived at a frame
	size_t arch_stage_cap = 0;
	// the records of the last push travel device -> pinned staging while the host goes on with the
	// next image; finishRecords appends them to `records` (their range is already in the journal)
	size_t pending_records = 0;
	cudaEvent_t rec_event = nullptr;  // recorded behind the copy (the stream may belong to another context)
	std::vector<ebt_opt_record> records;
	size_t records_sorted = 0;  // records[0 .. records_sorted) are in reference order
	std::vector<size_t> record_batches;  // end offset of every push behind records_sorted

	// pinned staging (device <-> host copies never touch pageable memory)
	DevPatch* h_stage = nullptr;
	ebt_opt_record* rec_stage = nullptr;
	size_t rec_stage_cap = 0;
	unsigned long long* nrec_stage = nullptr;
	std::vector<int> pos_of_id;  // track id -> index in h, rebuilt per push
	std::vector<int> load_est;   // per patch: expected optimisations in the next push (CTA order)

	// stream-kernel scratch: prefilter mask (P rows), CTA order
	unsigned* d_mask = nullptr;
	size_t mask_cap_words = 0;
	int* d_order = nullptr;
	int* h_order = nullptr;  // pinned
	double prefilter_margin = 8.0;

	ebt_event* d_events = nullptr;
	size_t d_events_cap = 0;
	unsigned char* d_gray = nullptr;
	double* d_lut = nullptr;

	ebt_problem* d_problems = nullptr;
	ebt_solution* d_solutions = nullptr;
	int* d_bnabla = nullptr;
	ebt_event* d_bevents = nullptr;
	size_t cap_problems = 0, cap_solutions = 0, cap_bnabla = 0, cap_bevents = 0;

	bool timing = false;
	cudaEvent_t ev0 = nullptr, ev1 = nullptr;
	// timing: event pairs are recorded without a host sync and resolved when the
	// totals are read (every launch is already followed by a stream sync of its own)
	struct Span { cudaEvent_t a, b; int k; };
	std::vector<Span> spans;
	std::vector<cudaEvent_t> event_pool;
	cudaEvent_t span_begin = nullptr;
	double ms[K_COUNT] = {};
	int64_t launches[K_COUNT] = {};

	Frontend fe;

	// event-text parser scratch (ebt_ingest.cu)
	char* d_text = nullptr;
	size_t text_cap = 0;
	int* d_tile_counts = nullptr;
	size_t tiles_cap = 0;
	long long* d_total = nullptr;          // + error word behind it
	long long* h_total = nullptr;          // pinned, 2 words
	ebt_event* d_parsed = nullptr;
	size_t parsed_cap = 0;
};

namespace
{
int fail(ebt_ctx* c, int code, const std::string& msg)
{
	if (c)
		c->err = msg;
	else
		g_create_error = msg;
	return code;
}

#define CU(call)                                                                          \
	do                                                                                    \
	{                                                                                     \
		cudaError_t e_ = (call);                                                          \
		if (e_ != cudaSuccess)                                                            \
			return fail(ctx, EBT_E_CUDA,                                                  \
						std::string(#call) + ": " + cudaGetErrorString(e_));             \
	} while (0)

// The kernels read the grid pointers from a device array; bring it up to date
// (new grids, a grown table) before a launch that uses it.
int syncFrameTable(ebt_ctx* ctx)
{
	if (!ctx->frame_ptrs_dirty) return EBT_OK;
	const size_t n = ctx->frames.size();
	if (n > ctx->frame_ptrs_cap)
	{
		CU(cudaStreamSynchronize(ctx->stream));  // kernels in flight read the old table
		cudaFree(ctx->d_frame_ptrs);
		ctx->d_frame_ptrs = nullptr;
		ctx->frame_ptrs_cap = 0;
		CU(cudaMalloc(&ctx->d_frame_ptrs, n * sizeof(const double2*)));
		ctx->frame_ptrs_cap = n;
	}
	ctx->h_frame_ptrs.resize(n);
	for (size_t i = 0; i < n; ++i) ctx->h_frame_ptrs[i] = ctx->frames[i].grid;
	// pageable source: the copy is staged before the call returns
	CU(cudaMemcpyAsync(ctx->d_frame_ptrs, ctx->h_frame_ptrs.data(), n * sizeof(const double2*), cudaMemcpyHostToDevice,
					   ctx->stream));
	ctx->frame_ptrs_dirty = false;
	return EBT_OK;
}

FrameTable frameTable(const ebt_ctx* c)
{
	FrameTable ft;
	ft.grid = c->d_frame_ptrs;
	ft.latest = c->latest;
	ft.pad = 0;
	return ft;
}

// A slot for a new template frame: a dead one, or a new one at the end of the
// table (FeatureDetector::optimizers_ is an unbounded map).
int allocFrameSlot(ebt_ctx* c)
{
	for (size_t i = 0; i < c->frames.size(); ++i)
		if (!c->frames[i].alive) return static_cast<int>(i);
	c->frames.resize(c->frames.size() * 2);
	c->frame_ptrs_dirty = true;
	return static_cast<int>(c->frames.size() / 2);
}

cudaEvent_t takeEvent(ebt_ctx* c)
{
	if (!c->event_pool.empty())
	{
		cudaEvent_t e = c->event_pool.back();
		c->event_pool.pop_back();
		return e;
	}
	cudaEvent_t e = nullptr;
	cudaEventCreate(&e);
	return e;
}
void resolveSpans(ebt_ctx* c)
{
	for (const auto& s : c->spans)
	{
		float ms = 0;
		if (cudaEventSynchronize(s.b) == cudaSuccess && cudaEventElapsedTime(&ms, s.a, s.b) == cudaSuccess)
			c->ms[s.k] += ms;
		c->event_pool.push_back(s.a);
		c->event_pool.push_back(s.b);
	}
	c->spans.clear();
}
void tic(ebt_ctx* c)
{
	if (!c->timing) return;
	c->span_begin = takeEvent(c);
	cudaEventRecord(c->span_begin, c->stream);
}
void toc(ebt_ctx* c, int k)
{
	c->launches[k]++;
	if (!c->timing || !c->span_begin) return;
	cudaEvent_t e = takeEvent(c);
	cudaEventRecord(e, c->stream);
	c->spans.push_back({c->span_begin, e, k});
	c->span_begin = nullptr;
	if (c->spans.size() >= 4096) resolveSpans(c);
}

void cornerOf(const DevPatch& p, double& x, double& y)
{
	x = p.rx + (EBT_PATCH_SIDE - 1) / 2.;
	y = p.ry + (EBT_PATCH_SIDE - 1) / 2.;
}

// Patch::addTrajectoryPosition at image time (feature_detector.cpp:497-506)
void addTraj(ebt_ctx* c, DevPatch& p)
{
	double x, y;
	cornerOf(p, x, y);
	c->img_samples.push_back({p.track_id, x, y, static_cast<double>(p.cur_ts)});
	p.n_traj++;
	p.traj_x[0] = p.traj_x[1];
	p.traj_y[0] = p.traj_y[1];
	p.traj_ts[0] = p.traj_ts[1];
	p.traj_x[1] = x;
	p.traj_y[1] = y;
	p.traj_ts[1] = p.cur_ts;
}

// Wait for the records of the last push and put them into their reserved place in c->records.
// Everything that reads c->records or reuses the staging buffer calls this first.
void finishRecords(ebt_ctx* c)
{
	if (c->pending_records == 0) return;
	cudaEventSynchronize(c->rec_event);
	// nothing else appends to c->records, so the range the journal already names starts at its end
	c->records.insert(c->records.end(), c->rec_stage, c->rec_stage + c->pending_records);
	c->pending_records = 0;
}

void materialise(ebt_ctx* c)
{
	finishRecords(c);
	if (c->all_tracks.size() < static_cast<size_t>(c->next_track_id))
		c->all_tracks.resize(static_cast<size_t>(c->next_track_id));
	for (const auto& e : c->journal)
	{
		if (e.kind == 0)
		{
			for (size_t i = e.a; i < e.b; ++i)
			{
				const auto& sm = c->img_samples[i];
				HostTrack& t = c->all_tracks[static_cast<size_t>(sm.id)];
				t.traj.push_back(sm.x);
				t.traj.push_back(sm.y);
				t.traj.push_back(sm.ts);
			}
		}
		else
		{
			// a feature's records are in the order it produced them, before and
			// after sortRecords (which orders by event first)
			for (size_t i = e.a; i < e.b; ++i)
			{
				const ebt_opt_record& r = c->records[i];
				HostTrack& t = c->all_tracks[static_cast<size_t>(r.track_id)];
				t.costs.push_back(r.final_cost);  // Patch::addFinalCost
				if (!(r.lost & 2))                // not dropped by the cost filter (optimizer.cpp:151-156)
				{
					t.traj.push_back(r.x);
					t.traj.push_back(r.y);
					t.traj.push_back(static_cast<double>(r.ts));
				}
			}
		}
	}
	c->journal.clear();
	c->img_samples.clear();
}

int findFrame(const ebt_ctx* c, int64_t ts)
{
	for (size_t i = 0; i < c->frames.size(); ++i)
		if (c->frames[i].alive && c->frames[i].ts == ts) return static_cast<int>(i);
	return -1;
}

// Upload one gray frame and build its gradient grid in `slot`.
int buildGrid(ebt_ctx* ctx, const uint8_t* gray, int slot, bool on_device = false)
{
	const size_t npx = static_cast<size_t>(ctx->prm.width) * ctx->prm.height;
	Frame& f = ctx->frames[slot];
	if (!f.grid)
	{
		CU(cudaMalloc(&f.grid, npx * sizeof(double2)));
		ctx->frame_ptrs_dirty = true;
	}
	CU(cudaMemcpyAsync(ctx->d_gray, gray, npx,
					   on_device ? cudaMemcpyDeviceToDevice : cudaMemcpyHostToDevice, ctx->stream));
	tic(ctx);
	CU(ebt_launch_grad(ctx->d_gray, ctx->d_lut, ctx->prm.width, ctx->prm.height, f.grid,
					   ctx->stream));
	toc(ctx, K_GRAD);
	// a host source may be reused by the caller as soon as we return
	if (!on_device) CU(cudaStreamSynchronize(ctx->stream));
	return EBT_OK;
}

int uploadPatches(ebt_ctx* ctx)
{
	if (!ctx->host_dirty || ctx->h.empty())
	{
		ctx->host_dirty = false;
		return EBT_OK;
	}
	// the previous use of h_stage (a download) was synchronised before h was read
	std::memcpy(ctx->h_stage, ctx->h.data(), ctx->h.size() * sizeof(DevPatch));
	CU(cudaMemcpyAsync(ctx->d_patches, ctx->h_stage, ctx->h.size() * sizeof(DevPatch),
					   cudaMemcpyHostToDevice, ctx->stream));
	ctx->host_dirty = false;
	return EBT_OK;
}

int downloadPatches(ebt_ctx* ctx)
{
	if (ctx->h.empty()) return EBT_OK;
	CU(cudaMemcpyAsync(ctx->h_stage, ctx->d_patches, ctx->h.size() * sizeof(DevPatch),
					   cudaMemcpyDeviceToHost, ctx->stream));
	CU(cudaStreamSynchronize(ctx->stream));
	std::memcpy(ctx->h.data(), ctx->h_stage, ctx->h.size() * sizeof(DevPatch));
	return EBT_OK;
}

// Reference order of optimisations: by event, then by position in patches_
// (feature_detector.cpp:585-619).  patches_ only ever appends new tracks (with
// increasing ids) and erases in place, so position order is track-id order.
void sortRecords(ebt_ctx* ctx)
{
	finishRecords(ctx);
	size_t lo = ctx->records_sorted;
	for (size_t hi : ctx->record_batches)
	{
		std::sort(ctx->records.begin() + static_cast<std::ptrdiff_t>(lo),
				  ctx->records.begin() + static_cast<std::ptrdiff_t>(hi),
				  [](const ebt_opt_record& a, const ebt_opt_record& b) {
					  if (a.event_index != b.event_index) return a.event_index < b.event_index;
					  return a.track_id < b.track_id;
				  });
		lo = hi;
	}
	ctx->records_sorted = ctx->records.size();
	ctx->record_batches.clear();
}

void fillTrack(const DevPatch& p, ebt_track& o)
{
	std::memset(&o, 0, sizeof(o));
	o.track_id = p.track_id;
	o.lost = p.lost;
	o.init = p.init;
	o.num_of_events = p.num_of_events;
	o.counter = p.counter;
	o.window_size = p.win_size;
	o.n_optimizations = p.n_costs;
	o.n_trajectory = p.n_traj;
	o.rect_x = p.rx;
	o.rect_y = p.ry;
	o.init_x = p.ix;
	o.init_y = p.iy;
	std::memcpy(o.warp, p.warp, sizeof(o.warp));
	o.flow_dir = p.init ? static_cast<double>(static_cast<float>(p.flow)) : 0.0;
	o.last_cost = p.n_costs > 0 ? p.costs[(p.n_costs - 1) & (EBT_COST_HISTORY - 1)] : 0.0;
	o.current_ts = p.cur_ts;
	o.time_last_update = p.time_last_update;
	o.time_without_update = p.time_without_update;
	o.init_time = p.init_time;
}
}  // namespace

extern "C" {

int ebt_default_params(ebt_params* p, int width, int height)
{
	if (!p || width <= 0 || height <= 0) return EBT_E_INVALID;
	std::memset(p, 0, sizeof(*p));
	p->width = width;
	p->height = height;
	p->max_patches = 100;
	p->init_num_events = 75;
	p->min_num_events = 100;
	p->max_num_events = 300;
	p->batch_stride = 30;
	p->max_iterations = 10;
	p->border_margin = 5;
	p->record_capacity = 1 << 18;
	p->association_distance = 5.0;
	p->huber_loss = 0.3;
	p->optimizer_threshold = 0.6;
	return EBT_OK;
}

const char* ebt_last_error(const ebt_ctx* ctx)
{
	return ctx ? ctx->err.c_str() : g_create_error.c_str();
}

int ebt_destroy(ebt_ctx* ctx)
{
	if (!ctx) return EBT_E_INVALID;
	cudaSetDevice(ctx->device);
	finishRecords(ctx);
	if (ctx->rec_event) cudaEventDestroy(ctx->rec_event);
	for (auto& f : ctx->frames) cudaFree(f.grid);
	cudaFree(ctx->d_frame_ptrs);
	cudaFree(ctx->d_patches);
	cudaFree(ctx->d_rings);
	cudaFree(ctx->d_nablas);
	cudaFree(ctx->d_mc_nablas);
	cudaFree(ctx->d_pred_nablas);
	cudaFree(ctx->d_cost_maps);
	cudaFree(ctx->d_bcost);
	cudaFree(ctx->d_records);
	cudaFree(ctx->d_nrec);
	cudaFree(ctx->d_queue);
	cudaFree(ctx->d_descs);
	cudaFreeHost(ctx->h_descs);
	cudaFree(ctx->d_work);
	cudaFreeHost(ctx->h_work);
	if (ctx->join_event) cudaEventDestroy(ctx->join_event);
	cudaFree(ctx->d_events);
	cudaFree(ctx->d_gray);
	cudaFree(ctx->d_lut);
	cudaFree(ctx->d_problems);
	cudaFree(ctx->d_solutions);
	cudaFree(ctx->d_bnabla);
	cudaFree(ctx->d_bevents);
	cudaFree(ctx->d_mask);
	cudaFree(ctx->d_order);
	cudaFreeHost(ctx->h_order);
	cudaFreeHost(ctx->h_stage);
	cudaFreeHost(ctx->arch_stage);
	for (int i = 0; i < 2; ++i)
	{
		cudaFree(ctx->d_replay[i]);
		cudaFreeHost(ctx->h_replay[i]);
		if (ctx->uploaded[i]) cudaEventDestroy(ctx->uploaded[i]);
		for (int j = 0; j < 2; ++j)
			if (ctx->chunk_done[i][j]) cudaEventDestroy(ctx->chunk_done[i][j]);
	}
	if (ctx->copy_stream) cudaStreamDestroy(ctx->copy_stream);
	cudaFreeHost(ctx->rec_stage);
	cudaFreeHost(ctx->nrec_stage);
	cudaFree(ctx->d_text);
	cudaFree(ctx->d_tile_counts);
	cudaFree(ctx->d_total);
	cudaFreeHost(ctx->h_total);
	cudaFree(ctx->d_parsed);
	{
		Frontend& fe = ctx->fe;
		for (int l = 0; l < EBT_LK_MAX_LEVELS; ++l)
		{
			cudaFree(fe.img[0][l]);
			cudaFree(fe.img[1][l]);
			cudaFree(fe.deriv[l]);
		}
		cudaFree(fe.scratch);
		cudaFree(fe.planes);
		cudaFree(fe.resp);
		cudaFree(fe.max_key);
		cudaFree(fe.n_cand);
		cudaFree(fe.cand);
		cudaFree(fe.d_pts);
		cudaFree(fe.d_status);
		cudaFree(fe.d_land);
		cudaFreeHost(fe.h_cand);
		cudaFreeHost(fe.h_ncand);
		cudaFreeHost(fe.h_pts);
		cudaFreeHost(fe.h_status);
	}
	resolveSpans(ctx);
	for (cudaEvent_t e : ctx->event_pool) cudaEventDestroy(e);
	if (ctx->ev0) cudaEventDestroy(ctx->ev0);
	if (ctx->ev1) cudaEventDestroy(ctx->ev1);
	if (ctx->stream) cudaStreamDestroy(ctx->stream);
	delete ctx;
	return EBT_OK;
}

int ebt_create(const ebt_params* params, int device, ebt_ctx** out)
{
	ebt_ctx* ctx = nullptr;  // errors before allocation go to g_create_error
	if (!params || !out) return fail(ctx, EBT_E_INVALID, "ebt_create: null argument");
	if (params->width <= 0 || params->height <= 0 || params->max_patches <= 0 ||
		params->max_num_events > EBT_MAX_WINDOW || params->min_num_events < 1 ||
		params->min_num_events > params->max_num_events || params->record_capacity <= 0)
		return fail(ctx, EBT_E_INVALID, "ebt_create: invalid parameters");
	int n_dev = 0;
	if (cudaGetDeviceCount(&n_dev) != cudaSuccess || n_dev <= 0)
		return fail(ctx, EBT_E_NODEVICE,
					"ebt_create: no CUDA device (this library has no CPU fallback)");
	if (device < 0 || device >= n_dev) return fail(ctx, EBT_E_INVALID, "ebt_create: bad device index");
	{
		cudaError_t e = cudaSetDevice(device);
		if (e != cudaSuccess)
			return fail(ctx, EBT_E_CUDA, std::string("cudaSetDevice: ") + cudaGetErrorString(e));
	}
	ebt_ctx* c = new ebt_ctx();
	c->prm = *params;
	c->device = device;
	ebt_default_frontend_params(&c->fe.prm);
	DevParams& d = c->dp;
	d.width = params->width;
	d.height = params->height;
	d.max_iterations = params->max_iterations;
	d.init_num_events = params->init_num_events;
	d.min_num_events = params->min_num_events;
	d.max_num_events = params->max_num_events;
	d.batch_stride = params->batch_stride;
	d.border_margin = params->border_margin;
	d.huber_a = params->huber_loss;
	d.optimizer_threshold = params->optimizer_threshold;
	// Ceres 2.0 Solver::Options defaults (ceres/include/ceres/solver.h:264-322)
	d.initial_radius = 1e4;
	d.max_radius = 1e16;
	d.min_radius = 1e-32;
	d.min_relative_decrease = 1e-3;
	d.min_lm_diagonal = 1e-6;
	d.max_lm_diagonal = 1e32;
	d.function_tolerance = 1e-6;
	d.gradient_tolerance = 1e-10;
	d.parameter_tolerance = 1e-8;
	d.max_consecutive_nonmonotonic_steps = 5;
	d.max_consecutive_invalid_steps = 5;
	d.products = 0;
	d.pad_ = 0;
	d.mc_nablas = nullptr;
	d.pred_nablas = nullptr;
	d.cost_maps = nullptr;

	ctx = c;
	const size_t P = static_cast<size_t>(params->max_patches);
	const size_t npx = static_cast<size_t>(params->width) * params->height;
#define CUC(call)                                                                         \
	do                                                                                    \
	{                                                                                     \
		cudaError_t e_ = (call);                                                          \
		if (e_ != cudaSuccess)                                                            \
		{                                                                                 \
			g_create_error = std::string(#call) + ": " + cudaGetErrorString(e_);         \
			ebt_destroy(c);                                                               \
			return EBT_E_CUDA;                                                            \
		}                                                                                 \
	} while (0)
	CUC(cudaDeviceGetAttribute(&c->sm_count, cudaDevAttrMultiProcessorCount, device));
	CUC(cudaStreamCreateWithFlags(&c->stream, cudaStreamNonBlocking));
	CUC(cudaEventCreate(&c->ev0));
	CUC(cudaEventCreate(&c->ev1));
	CUC(cudaMalloc(&c->d_patches, P * sizeof(DevPatch)));
	CUC(cudaMalloc(&c->d_rings, P * EBT_RING * sizeof(ebt_event)));
	CUC(cudaMalloc(&c->d_nablas, P * EBT_PATCH_PIXELS * sizeof(int)));
	CUC(cudaMemset(c->d_nablas, 0, P * EBT_PATCH_PIXELS * sizeof(int)));
	CUC(cudaMalloc(&c->d_records, static_cast<size_t>(params->record_capacity) * sizeof(ebt_opt_record)));
	CUC(cudaMalloc(&c->d_nrec, sizeof(unsigned long long)));
	CUC(cudaMalloc(&c->d_queue, sizeof(int)));
	CUC(cudaHostAlloc(&c->h_stage, P * sizeof(DevPatch), cudaHostAllocDefault));
	CUC(cudaMalloc(&c->d_order, P * sizeof(int)));
	CUC(cudaHostAlloc(&c->h_order, P * sizeof(int), cudaHostAllocDefault));
	CUC(cudaHostAlloc(&c->nrec_stage, sizeof(unsigned long long), cudaHostAllocDefault));
	CUC(cudaMalloc(&c->d_gray, npx));
	CUC(cudaMalloc(&c->d_lut, 256 * sizeof(double)));
	{
		// FeatureDetector::getLogImage (feature_detector.cpp:714-723): the input
		// is 8-bit, so log(I/255 + 0.1)/8 is a 256-entry table, computed once on
		// the host with the same libm the reference uses.
		double lut[256];
		for (int i = 0; i < 256; ++i)
		{
			volatile double v = i * (1.0 / 255.0);  // convertTo, then "+ 10e-2": two roundings
			v = v + 10e-2;
			lut[i] = std::log(v) * (1.0 / 8);
		}
		CUC(cudaMemcpy(c->d_lut, lut, sizeof(lut), cudaMemcpyHostToDevice));
	}
#undef CUC
	for (int s = static_cast<int>(P) - 1; s >= 0; --s) c->free_slots.push_back(s);
	*out = c;
	return EBT_OK;
}

// Back to the state right after ebt_create (FeatureDetector::setParams ->
// reset(), feature_detector.cpp:734-742), keeping all allocations.
int ebt_reset(ebt_ctx* ctx)
{
	if (!ctx) return EBT_E_INVALID;
	CU(cudaSetDevice(ctx->device));
	CU(cudaStreamSynchronize(ctx->stream));
	for (auto& f : ctx->frames)
	{
		f.alive = false;
		f.users = 0;
	}
	ctx->latest = 0;
	ctx->h.clear();
	ctx->all_tracks.clear();
	ctx->img_samples.clear();
	ctx->journal.clear();
	ctx->arch.clear();
	ctx->arch_nabla.clear();
	ctx->arch_mc.clear();
	ctx->arch_pred.clear();
	ctx->arch_cost.clear();
	finishRecords(ctx);
	ctx->records.clear();
	ctx->records_sorted = 0;
	ctx->record_batches.clear();
	ctx->free_slots.clear();
	for (int s = ctx->prm.max_patches - 1; s >= 0; --s) ctx->free_slots.push_back(s);
	ctx->host_dirty = true;
	ctx->next_track_id = 0;
	ctx->image_counter = 0;
	ctx->event_index = 0;
	ctx->corners.clear();
	ctx->fe.have[0] = ctx->fe.have[1] = false;
	ctx->fe.built[0] = ctx->fe.built[1] = 0;
	CU(cudaMemsetAsync(ctx->d_nablas, 0,
					   static_cast<size_t>(ctx->prm.max_patches) * EBT_PATCH_PIXELS * sizeof(int), ctx->stream));
	if (ctx->d_cost_maps)
		CU(cudaMemsetAsync(ctx->d_cost_maps, 0, static_cast<size_t>(ctx->prm.max_patches) * 121 * sizeof(double),
						   ctx->stream));
	if (ctx->d_mc_nablas)
	{
		CU(cudaMemsetAsync(ctx->d_mc_nablas, 0,
						   static_cast<size_t>(ctx->prm.max_patches) * EBT_PATCH_PIXELS * sizeof(int), ctx->stream));
		CU(cudaMemsetAsync(ctx->d_pred_nablas, 0,
						   static_cast<size_t>(ctx->prm.max_patches) * EBT_PATCH_PIXELS * sizeof(double), ctx->stream));
	}
	return EBT_OK;
}

// Measured non-tensor FMA throughput (TFLOP/s) of this device: the roofline
// denominator for the LM kernels (SURVEY.md 8d asks for a measured figure).
int ebt_measure_fma_peak(ebt_ctx* ctx, double* fp64_tflops, double* fp32_tflops)
{
	if (!ctx) return EBT_E_INVALID;
	CU(cudaSetDevice(ctx->device));
	cudaDeviceProp prop;
	CU(cudaGetDeviceProperties(&prop, ctx->device));
	const int blocks = prop.multiProcessorCount * 8, threads = 256, iters = 4096;
	void* buf = nullptr;
	CU(cudaMalloc(&buf, static_cast<size_t>(blocks) * threads * sizeof(double)));
	for (int fp64 = 1; fp64 >= 0; --fp64)
	{
		double best = 0;
		for (int rep = 0; rep < 5; ++rep)
		{
			CU(cudaEventRecord(ctx->ev0, ctx->stream));
			CU(ebt_launch_fma_peak(fp64, buf, blocks, threads, iters, ctx->stream));
			CU(cudaEventRecord(ctx->ev1, ctx->stream));
			CU(cudaEventSynchronize(ctx->ev1));
			float ms = 0;
			CU(cudaEventElapsedTime(&ms, ctx->ev0, ctx->ev1));
			const double tf = 2.0 * 8 * iters * static_cast<double>(blocks) * threads / (ms * 1e-3) / 1e12;
			if (rep > 0 && tf > best) best = tf;
		}
		if (fp64 && fp64_tflops) *fp64_tflops = best;
		if (!fp64 && fp32_tflops) *fp32_tflops = best;
	}
	cudaFree(buf);
	return EBT_OK;
}

int ebt_sync(ebt_ctx* ctx)
{
	if (!ctx) return EBT_E_INVALID;
	CU(cudaSetDevice(ctx->device));
	CU(cudaStreamSynchronize(ctx->stream));
	return EBT_OK;
}

int ebt_timing_enable(ebt_ctx* ctx, int on)
{
	if (!ctx) return EBT_E_INVALID;
	ctx->timing = on != 0;
	return EBT_OK;
}
int ebt_timing_reset(ebt_ctx* ctx)
{
	if (!ctx) return EBT_E_INVALID;
	resolveSpans(ctx);
	for (int k = 0; k < K_COUNT; ++k)
	{
		ctx->ms[k] = 0;
		ctx->launches[k] = 0;
	}
	return EBT_OK;
}
int ebt_timing_get(ebt_ctx* ctx, int kernel, double* ms_total, int64_t* launches)
{
	if (!ctx || kernel < 0 || kernel >= K_COUNT) return EBT_E_INVALID;
	resolveSpans(ctx);
	if (ms_total) *ms_total = ctx->ms[kernel];
	if (launches) *launches = ctx->launches[kernel];
	return EBT_OK;
}

size_t ebt_sizeof_event(void) { return sizeof(ebt_event); }
size_t ebt_sizeof_track(void) { return sizeof(ebt_track); }
size_t ebt_sizeof_opt_record(void) { return sizeof(ebt_opt_record); }
size_t ebt_sizeof_problem(void) { return sizeof(ebt_problem); }
size_t ebt_sizeof_solution(void) { return sizeof(ebt_solution); }
size_t ebt_sizeof_params(void) { return sizeof(ebt_params); }
size_t ebt_sizeof_frontend_params(void) { return sizeof(ebt_frontend_params); }
size_t ebt_sizeof_camera(void) { return sizeof(ebt_camera); }
int ebt_abi_version(void) { return 1; }

int ebt_debug_phase_clocks(ebt_ctx* ctx, uint64_t* out8)
{
	if (!ctx || !out8) return fail(ctx, EBT_E_INVALID, "ebt_debug_phase_clocks: bad argument");
	CU(cudaSetDevice(ctx->device));
	unsigned long long v[8];
	CU(ebt_launch_phase_clocks(v, ctx->stream));
	for (int i = 0; i < 8; ++i) out8[i] = v[i];
	return EBT_OK;
}

int ebt_debug_feature_load(ebt_ctx* ctx, int32_t* track_ids, int32_t* n_opt, uint64_t* t0_ns, uint64_t* t1_ns,
						   int32_t* sm, int cap, int* n_out)
{
	if (!ctx || !n_out || cap < 0) return fail(ctx, EBT_E_INVALID, "ebt_debug_feature_load: bad argument");
	int n = 0;
	for (const DevPatch& p : ctx->h)
	{
		if (n < cap)
		{
			if (track_ids) track_ids[n] = p.track_id;
			if (n_opt) n_opt[n] = p.last_n_opt;
			if (t0_ns) t0_ns[n] = p.last_t0_ns;
			if (t1_ns) t1_ns[n] = p.last_t1_ns;
			if (sm) sm[n] = p.last_sm;
		}
		++n;
	}
	*n_out = n;
	return EBT_OK;
}

int ebt_set_stream_kernel(ebt_ctx* ctx, int mode)
{
	if (!ctx || mode < 0 || mode > 2) return fail(ctx, EBT_E_INVALID, "ebt_set_stream_kernel: mode must be 0, 1 or 2");
	ctx->stream_kernel = mode;
	return EBT_OK;
}

int ebt_set_prefilter_margin(ebt_ctx* ctx, double pixels)
{
	if (!ctx || !(pixels >= 0.0) || pixels > 4096.0)
		return fail(ctx, EBT_E_INVALID, "ebt_set_prefilter_margin: margin must be in [0, 4096] px");
	ctx->prefilter_margin = pixels;
	return EBT_OK;
}

// FeatureDetector::newImage, first half: extractPatches (feature_detector.cpp:
// 543-566) with injected corners, associatePatches (:630-664).
static int newFrameBegin(ebt_ctx* ctx, const uint8_t* gray, int64_t ts_us, const double* cornersXY,
						 int n_corners, bool on_device);
int ebt_new_frame_begin(ebt_ctx* ctx, const uint8_t* gray, int64_t ts_us, const double* cornersXY,
						int n_corners)
{
	return newFrameBegin(ctx, gray, ts_us, cornersXY, n_corners, false);
}
int ebt_new_frame_begin_device(ebt_ctx* ctx, const uint8_t* d_gray, int64_t ts_us,
							   const double* cornersXY, int n_corners)
{
	return newFrameBegin(ctx, d_gray, ts_us, cornersXY, n_corners, true);
}
static int newFrameBegin(ebt_ctx* ctx, const uint8_t* gray, int64_t ts_us, const double* cornersXY,
						 int n_corners, bool on_device)
{
	if (!ctx || !gray || n_corners < 0 || (n_corners > 0 && !cornersXY))
		return fail(ctx, EBT_E_INVALID, "ebt_new_frame_begin: bad argument");
	CU(cudaSetDevice(ctx->device));
	int slot = findFrame(ctx, ts_us);
	if (slot < 0) slot = allocFrameSlot(ctx);
	{
		const int rc = buildGrid(ctx, gray, slot, on_device);
		if (rc != EBT_OK) return rc;
	}
	Frame& f = ctx->frames[slot];
	f.ts = ts_us;
	f.users = 0;
	f.alive = true;
	ctx->latest = slot;

	if (n_corners > 0 && cornersXY != ctx->corners.data())
		ctx->corners.assign(cornersXY, cornersXY + 2 * static_cast<size_t>(n_corners));
	else if (n_corners == 0)
		ctx->corners.clear();
	std::vector<int> newIds(static_cast<size_t>(n_corners), -1);
	const size_t first_sample = ctx->img_samples.size();
	// associatePatches (feature_detector.cpp:630-664): every patch, in list
	// order, takes the FIRST corner (in corner order) that is still free and
	// closer than associationDistance.  The corners are bucketed on a uniform
	// grid of cell size d, so a patch only looks at the 3x3 cells around its
	// centre; each bucket keeps corner order, so "first" is the smallest index
	// over the nine buckets — the same decisions as the reference's double loop.
	const double d = ctx->prm.association_distance;
	bool use_grid = d > 0.0 && static_cast<size_t>(n_corners) * ctx->h.size() > 4096;
	for (int j = 0; j < 2 * n_corners && use_grid; ++j) use_grid = std::isfinite(cornersXY[j]);
	const int gw = use_grid ? static_cast<int>(std::min(4096.0, std::ceil(ctx->prm.width / d) + 2.0)) : 0;
	const int gh = use_grid ? static_cast<int>(std::min(4096.0, std::ceil(ctx->prm.height / d) + 2.0)) : 0;
	auto cellOf = [&](double v, int n) {
		const double c = std::floor(v / d) + 1.0;  // monotone in v, then clamped: neighbours stay neighbours
		return c < 0.0 ? 0 : (c > n - 1 ? n - 1 : static_cast<int>(c));
	};
	std::vector<int>& cellStart = ctx->assoc_start;
	std::vector<int>& cellItems = ctx->assoc_items;
	if (use_grid)
	{
		cellStart.assign(static_cast<size_t>(gw) * gh + 1, 0);
		cellItems.resize(static_cast<size_t>(n_corners));
		for (int j = 0; j < n_corners; ++j)
			cellStart[static_cast<size_t>(cellOf(cornersXY[2 * j + 1], gh)) * gw + cellOf(cornersXY[2 * j], gw) + 1]++;
		for (size_t c = 1; c < cellStart.size(); ++c) cellStart[c] += cellStart[c - 1];
		std::vector<int>& fill = ctx->assoc_fill;
		fill.assign(cellStart.begin(), cellStart.end() - 1);
		for (int j = 0; j < n_corners; ++j)
			cellItems[static_cast<size_t>(
				fill[static_cast<size_t>(cellOf(cornersXY[2 * j + 1], gh)) * gw + cellOf(cornersXY[2 * j], gw)]++)] = j;
	}
	auto matches = [&](double cx, double cy, int j) {
		const double dx = cx - cornersXY[2 * j], dy = cy - cornersXY[2 * j + 1];
		// |dx| or |dy| >= d implies the norm is >= d: skip the sqrt, same decision
		if (!(std::fabs(dx) < d) || !(std::fabs(dy) < d)) return false;
		return newIds[static_cast<size_t>(j)] == -1 && std::sqrt(dx * dx + dy * dy) < d;
	};
	for (size_t i = 0; i < ctx->h.size(); ++i)
	{
		DevPatch& p = ctx->h[i];
		double cx, cy;
		cornerOf(p, cx, cy);
		if (use_grid && std::isfinite(cx) && std::isfinite(cy))
		{
			int best = n_corners;
			const int gx = cellOf(cx, gw), gy = cellOf(cy, gh);
			for (int yy = std::max(gy - 1, 0); yy <= std::min(gy + 1, gh - 1); ++yy)
				for (int xx = std::max(gx - 1, 0); xx <= std::min(gx + 1, gw - 1); ++xx)
				{
					const size_t c = static_cast<size_t>(yy) * gw + xx;
					for (int k = cellStart[c]; k < cellStart[c + 1]; ++k)
					{
						const int j = cellItems[static_cast<size_t>(k)];
						if (j >= best) break;  // buckets are in corner order
						if (matches(cx, cy, j))
						{
							best = j;
							break;
						}
					}
				}
			if (best < n_corners) newIds[static_cast<size_t>(best)] = p.track_id;
		}
		else
			for (int j = 0; j < n_corners; ++j)
				if (matches(cx, cy, j))
				{
					newIds[static_cast<size_t>(j)] = p.track_id;
					break;
				}
		p.cur_ts = ts_us;  // Patch::setTs
		addTraj(ctx, p);
	}
	for (int j = 0; j < n_corners; ++j)
	{
		if (newIds[j] != -1 || static_cast<int>(ctx->h.size()) >= ctx->prm.max_patches) continue;
		// Patch::Patch + init (patch.cpp:8-35)
		DevPatch p;
		std::memset(&p, 0, sizeof(p));
		p.rx = cornersXY[2 * j] - EBT_PATCH_EXTENT;
		p.ry = cornersXY[2 * j + 1] - EBT_PATCH_EXTENT;
		cornerOf(p, p.ix, p.iy);
		p.warp[0] = 1.0;
		p.cur_ts = ts_us;
		p.time_without_update = static_cast<long long>(1e7);
		p.init_time = ts_us;
		p.track_id = ctx->next_track_id++;
		p.num_of_events = 75;  // patch.cpp:22
		p.frame_slot = slot;
		p.foreign = (p.track_id % ctx->shard_count) != ctx->shard_index;
		p.store_slot = ctx->free_slots.back();
		ctx->free_slots.pop_back();
		addTraj(ctx, p);
		ctx->h.push_back(p);
		f.users++;
	}
	if (ctx->img_samples.size() > first_sample)
		ctx->journal.push_back({0, first_sample, ctx->img_samples.size()});
	if (ctx->image_counter < 2) ctx->image_counter++;  // FlowEstimator::addImage
	ctx->host_dirty = true;
	return EBT_OK;
}

// FlowEstimator::getFlowPatches, query side (flow_estimator.cpp:27-39).
int ebt_get_uninit(ebt_ctx* ctx, int32_t* track_ids, double* cornersXY, int cap, int* n_out)
{
	if (!ctx || !n_out) return fail(ctx, EBT_E_INVALID, "ebt_get_uninit: bad argument");
	int n = 0;
	if (ctx->image_counter == 2)
		for (const DevPatch& p : ctx->h)
		{
			if (p.init) continue;
			if (n < cap && track_ids && cornersXY)
			{
				track_ids[n] = p.track_id;
				cornerOf(p, cornersXY[2 * n], cornersXY[2 * n + 1]);
			}
			++n;
		}
	*n_out = n;
	return EBT_OK;
}

// Loop body of FlowEstimator::getFlowPatches (flow_estimator.cpp:41-79).
static void applyFlow(ebt_ctx* ctx, DevPatch& p, int found, float next_x, float next_y)
{
	ctx->host_dirty = true;
	if (!found)
	{
		p.lost = 1;
		return;
	}
	double cx, cy;
	cornerOf(p, cx, cy);
	const double dx = next_x - cx, dy = next_y - cy;
	p.warp[0] = 1.0;
	p.warp[1] = 0.0;
	p.warp[2] = -dx;
	p.warp[3] = -dy;
	p.flow = std::atan2(dy, dx);
	p.init = 1;
	p.time_without_update = static_cast<long long>(1e6 / std::fmax(1e-1, std::sqrt(dx * dx + dy * dy)));
	if (cx <= 5 || cy <= 5 || cx >= ctx->prm.width - 5 || cy >= ctx->prm.height - 5) p.lost = 1;
}

int ebt_set_flow(ebt_ctx* ctx, int32_t track_id, int found, float next_x, float next_y)
{
	if (!ctx) return EBT_E_INVALID;
	for (DevPatch& p : ctx->h)
	{
		if (p.track_id != track_id || p.init) continue;
		applyFlow(ctx, p, found, next_x, next_y);
		return EBT_OK;
	}
	return fail(ctx, EBT_E_INVALID, "ebt_set_flow: no uninitialised patch with this track id");
}

// The same for n patches at once (what a caller that runs LK on all the
// uninitialised patches of a frame has in hand): nextXY holds n (x, y) pairs.
int ebt_set_flows(ebt_ctx* ctx, const int32_t* track_ids, const int32_t* found, const float* nextXY, int n)
{
	if (!ctx || n < 0 || (n > 0 && (!track_ids || !found || !nextXY)))
		return fail(ctx, EBT_E_INVALID, "ebt_set_flows: bad argument");
	ctx->pos_of_id.assign(static_cast<size_t>(ctx->next_track_id), -1);
	for (size_t i = 0; i < ctx->h.size(); ++i)
		ctx->pos_of_id[static_cast<size_t>(ctx->h[i].track_id)] = static_cast<int>(i);
	for (int i = 0; i < n; ++i)
	{
		const int32_t id = track_ids[i];
		const int pos = (id >= 0 && id < ctx->next_track_id) ? ctx->pos_of_id[static_cast<size_t>(id)] : -1;
		if (pos < 0 || ctx->h[static_cast<size_t>(pos)].init)
			return fail(ctx, EBT_E_INVALID, "ebt_set_flows: no uninitialised patch with this track id");
		applyFlow(ctx, ctx->h[static_cast<size_t>(pos)], found[i], nextXY[2 * i], nextXY[2 * i + 1]);
	}
	return EBT_OK;
}

// FeatureDetector::newImage, second half (feature_detector.cpp:499-535).
// first half of newImage's tail: upload, updateNumOfEvents for every patch, state on its way back (queued only)
static int frameEndEnqueue(ebt_ctx* ctx)
{
	CU(cudaSetDevice(ctx->device));
	if (ctx->h.empty()) return EBT_OK;
	const int rc = uploadPatches(ctx);
	if (rc != EBT_OK) return rc;
	const int rcf = syncFrameTable(ctx);
	if (rcf != EBT_OK) return rcf;
	tic(ctx);
	CU(ebt_launch_num_events(ctx->d_patches, static_cast<int>(ctx->h.size()), frameTable(ctx), ctx->dp, ctx->stream));
	toc(ctx, K_NUMEV);
	CU(cudaMemcpyAsync(ctx->h_stage, ctx->d_patches, ctx->h.size() * sizeof(DevPatch), cudaMemcpyDeviceToHost, ctx->stream));
	return EBT_OK;
}

static int frameEndFinish(ebt_ctx* ctx);

int ebt_new_frame_end(ebt_ctx* ctx)
{
	if (!ctx) return EBT_E_INVALID;
	const int rc = frameEndEnqueue(ctx);
	return rc != EBT_OK ? rc : frameEndFinish(ctx);
}

// The same for the contexts of several streams (configs[3]): the device work of all of them is queued
// first, each on its own CUDA stream, then waited for — one host thread, the latencies overlap.
int ebt_new_frame_end_multi(ebt_ctx* const* ctxs, int n_streams)
{
	ebt_ctx* ctx = (ctxs && n_streams > 0) ? ctxs[0] : nullptr;
	if (!ctx) return EBT_E_INVALID;
	for (int i = 0; i < n_streams; ++i)
		if (!ctxs[i]) return fail(ctx, EBT_E_INVALID, "ebt_new_frame_end_multi: null context");
	int result = EBT_OK, queued = 0;
	for (; queued < n_streams && result == EBT_OK; ++queued) result = frameEndEnqueue(ctxs[queued]);
	if (result != EBT_OK) --queued;  // that one failed part-way: its staging buffer holds nothing to take back
	// every stream whose work was queued is finished, also after a failure elsewhere; the streams
	// behind a failed one were not touched and keep their state for a retry
	for (int i = 0; i < queued; ++i)
	{
		const int rc = frameEndFinish(ctxs[i]);
		if (rc != EBT_OK && result == EBT_OK) result = rc;
	}
	if (result != EBT_OK)
		for (int i = 1; i < n_streams; ++i)
			if (!ctxs[i]->err.empty() && ctx->err.empty()) ctx->err = ctxs[i]->err;
	return result;
}

static int frameEndFinish(ebt_ctx* ctx)
{
	CU(cudaSetDevice(ctx->device));
	if (!ctx->h.empty())
	{
		CU(cudaStreamSynchronize(ctx->stream));
		std::memcpy(ctx->h.data(), ctx->h_stage, ctx->h.size() * sizeof(DevPatch));
	}
	// archive lost patches, keep list order (feature_detector.cpp:512-523): their
	// patch products come back through one pinned staging buffer — every copy and
	// every clearing of a freed slot is queued on the context's stream, one wait
	size_t n_lost = 0;
	for (const DevPatch& p : ctx->h) n_lost += p.lost ? 1 : 0;
	if (n_lost > 0)
	{
		const bool prod = ctx->dp.products != 0, cmap = ctx->dp.cost_maps != nullptr;
		const size_t per = EBT_PATCH_PIXELS * sizeof(int) +
						   (prod ? EBT_PATCH_PIXELS * (sizeof(int) + sizeof(double)) : 0) + (cmap ? 121 * sizeof(double) : 0);
		if (n_lost * per > ctx->arch_stage_cap)
		{
			cudaFreeHost(ctx->arch_stage);
			ctx->arch_stage = nullptr;
			ctx->arch_stage_cap = 0;
			const size_t cap = n_lost * per * 2 + 65536;
			CU(cudaHostAlloc(&ctx->arch_stage, cap, cudaHostAllocDefault));
			ctx->arch_stage_cap = cap;
		}
		unsigned char* dst = ctx->arch_stage;
		for (const DevPatch& p : ctx->h)
		{
			if (!p.lost) continue;
			const size_t off = static_cast<size_t>(p.store_slot) * EBT_PATCH_PIXELS;
			CU(cudaMemcpyAsync(dst, ctx->d_nablas + off, EBT_PATCH_PIXELS * sizeof(int), cudaMemcpyDeviceToHost, ctx->stream));
			CU(cudaMemsetAsync(ctx->d_nablas + off, 0, EBT_PATCH_PIXELS * sizeof(int), ctx->stream));
			dst += EBT_PATCH_PIXELS * sizeof(int);
			if (prod)
			{
				CU(cudaMemcpyAsync(dst, ctx->d_mc_nablas + off, EBT_PATCH_PIXELS * sizeof(int), cudaMemcpyDeviceToHost, ctx->stream));
				CU(cudaMemsetAsync(ctx->d_mc_nablas + off, 0, EBT_PATCH_PIXELS * sizeof(int), ctx->stream));
				dst += EBT_PATCH_PIXELS * sizeof(int);
				CU(cudaMemcpyAsync(dst, ctx->d_pred_nablas + off, EBT_PATCH_PIXELS * sizeof(double), cudaMemcpyDeviceToHost, ctx->stream));
				CU(cudaMemsetAsync(ctx->d_pred_nablas + off, 0, EBT_PATCH_PIXELS * sizeof(double), ctx->stream));
				dst += EBT_PATCH_PIXELS * sizeof(double);
			}
			if (cmap)
			{
				const size_t coff = static_cast<size_t>(p.store_slot) * 121;
				CU(cudaMemcpyAsync(dst, ctx->d_cost_maps + coff, 121 * sizeof(double), cudaMemcpyDeviceToHost, ctx->stream));
				CU(cudaMemsetAsync(ctx->d_cost_maps + coff, 0, 121 * sizeof(double), ctx->stream));
				dst += 121 * sizeof(double);
			}
		}
		CU(cudaStreamSynchronize(ctx->stream));
		const unsigned char* src = ctx->arch_stage;
		size_t k = 0;
		for (size_t i = 0; i < ctx->h.size(); ++i)
		{
			const DevPatch& p = ctx->h[i];
			if (!p.lost)
			{
				if (k != i) ctx->h[k] = ctx->h[i];
				++k;
				continue;
			}
			const int* nab = reinterpret_cast<const int*>(src);
			ctx->arch_nabla.emplace_back(nab, nab + EBT_PATCH_PIXELS);
			src += EBT_PATCH_PIXELS * sizeof(int);
			if (prod)
			{
				const int* mc = reinterpret_cast<const int*>(src);
				ctx->arch_mc.emplace_back(mc, mc + EBT_PATCH_PIXELS);
				src += EBT_PATCH_PIXELS * sizeof(int);
				std::vector<double> pr(EBT_PATCH_PIXELS);
				std::memcpy(pr.data(), src, EBT_PATCH_PIXELS * sizeof(double));  // staged behind ints: not 8-byte aligned
				ctx->arch_pred.push_back(std::move(pr));
				src += EBT_PATCH_PIXELS * sizeof(double);
			}
			else
			{
				ctx->arch_mc.emplace_back();
				ctx->arch_pred.emplace_back();
			}
			if (cmap)
			{
				std::vector<double> cm(121);
				std::memcpy(cm.data(), src, 121 * sizeof(double));
				ctx->arch_cost.push_back(std::move(cm));
				src += 121 * sizeof(double);
			}
			else
				ctx->arch_cost.emplace_back();
			ctx->arch.push_back(p);
			ctx->free_slots.push_back(p.store_slot);
			ctx->frames[static_cast<size_t>(p.frame_slot)].users--;
		}
		ctx->h.resize(k);
		ctx->host_dirty = true;
	}
	// erase unused optimizers (feature_detector.cpp:525-535)
	for (auto& f : ctx->frames)
		if (f.alive && f.users == 0) f.alive = false;
	return EBT_OK;
}

// ---- one push = prefilter + stream kernel over one piece of the batch ---------
// Mask scratch of a context for a piece of n events.
static int ensureMask(ebt_ctx* ctx, int64_t n, int* wpr_out)
{
	const int wpr = ebt_mask_words_per_row(n);
	const size_t need_words = ctx->h.size() * static_cast<size_t>(wpr);
	if (need_words > ctx->mask_cap_words)
	{
		CU(cudaStreamSynchronize(ctx->stream));
		cudaFree(ctx->d_mask);
		ctx->d_mask = nullptr;
		ctx->mask_cap_words = 0;
		const size_t cap = need_words + need_words / 4;
		CU(cudaMalloc(&ctx->d_mask, cap * sizeof(unsigned)));
		ctx->mask_cap_words = cap;
	}
	*wpr_out = wpr;
	return EBT_OK;
}

// Expected optimisations per patch in the next push (work order: longest-
// processing-time-first keeps the kernel's tail short) = the previous push's
// count; a patch that has its flow but has never been optimised (created at the
// previous image, so its previous push only collected events) is expected to
// carry the mean load — rating it 0 would start a full-weight chain last.
static void estimateLoad(ebt_ctx* ctx)
{
	const int P = static_cast<int>(ctx->h.size());
	long long sum = 0, cnt = 0;
	for (int i = 0; i < P; ++i)
	{
		const DevPatch& p = ctx->h[static_cast<size_t>(i)];
		if (!p.lost && !p.foreign && p.init && p.n_costs > 0)
		{
			sum += p.last_n_opt;
			++cnt;
		}
	}
	const int mean = cnt ? static_cast<int>((sum + cnt - 1) / cnt) : 1;
	ctx->load_est.resize(static_cast<size_t>(P));
	for (int i = 0; i < P; ++i)
	{
		const DevPatch& p = ctx->h[static_cast<size_t>(i)];
		ctx->load_est[static_cast<size_t>(i)] =
			(p.lost || p.foreign) ? -1 : (!p.init ? 0 : (p.n_costs == 0 ? mean : p.last_n_opt));
	}
}

// After the kernel: patch state and record count back (one sync), then the
// records themselves (second sync).  `on` = the stream the kernel ran on.
static int collectPush(ebt_ctx* ctx, int64_t n, cudaStream_t on, int64_t* n_opt, bool synced_state)
{
	if (!synced_state)
	{
		CU(cudaMemcpyAsync(ctx->nrec_stage, ctx->d_nrec, sizeof(unsigned long long), cudaMemcpyDeviceToHost, on));
		CU(cudaMemcpyAsync(ctx->h_stage, ctx->d_patches, ctx->h.size() * sizeof(DevPatch), cudaMemcpyDeviceToHost, on));
		CU(cudaStreamSynchronize(on));
	}
	std::memcpy(ctx->h.data(), ctx->h_stage, ctx->h.size() * sizeof(DevPatch));
	const unsigned long long nrec = *ctx->nrec_stage;
	ctx->event_index += n;
	if (n_opt) *n_opt += static_cast<int64_t>(nrec);
	ctx->opt_per_event = std::max(static_cast<double>(nrec) / static_cast<double>(std::max<int64_t>(n, 1)), 1e-3);
	const unsigned long long got =
		std::min<unsigned long long>(nrec, static_cast<unsigned long long>(ctx->prm.record_capacity));
	if (got > 0)
	{
		if (got > ctx->rec_stage_cap)
		{
			cudaFreeHost(ctx->rec_stage);
			ctx->rec_stage = nullptr;
			ctx->rec_stage_cap = 0;
			const size_t cap = static_cast<size_t>(got) + static_cast<size_t>(got) / 2 + 4096;
			CU(cudaHostAlloc(&ctx->rec_stage, cap * sizeof(ebt_opt_record), cudaHostAllocDefault));
			ctx->rec_stage_cap = cap;
		}
		// The copy is queued and NOT waited for: it travels while the host handles the next image
		// (finishRecords).  Its place in the record list and in the journal is reserved now, so the
		// order of trajectory entries (image samples vs optimisations) stays the reference's.
		// A feature's records sit in the buffer in the order it produced them, which is all the
		// per-track histories need (built lazily, materialise()); the global reference order is
		// restored lazily too (sortRecords).
		CU(cudaMemcpyAsync(ctx->rec_stage, ctx->d_records, got * sizeof(ebt_opt_record), cudaMemcpyDeviceToHost, on));
		const size_t first = ctx->records.size();  // the previous push's records are in by now
		ctx->record_batches.push_back(first + static_cast<size_t>(got));
		ctx->journal.push_back({1, first, first + static_cast<size_t>(got)});
		if (!ctx->rec_event) CU(cudaEventCreateWithFlags(&ctx->rec_event, cudaEventDisableTiming));
		CU(cudaEventRecord(ctx->rec_event, on));
		ctx->pending_records = static_cast<size_t>(got);
	}
	if (nrec > got)
		return fail(ctx, EBT_E_CAPACITY,
					"ebt_push_events: record_capacity exceeded; tracks are valid, records/trajectories of this push are "
					"truncated (raise ebt_params::record_capacity)");
	return EBT_OK;
}

// One piece on the one-CTA-per-feature kernel (any feature count, products and cost map allowed).
static int runStream(ebt_ctx* ctx, const ebt_event* d_events, int64_t n, int64_t* n_opt)
{
	const int P = static_cast<int>(ctx->h.size());
	// the kernel below rewrites the device record buffer: it must start behind the previous push's record copy
	if (ctx->pending_records) CU(cudaStreamWaitEvent(ctx->stream, ctx->rec_event, 0));
	{
		const int rc = uploadPatches(ctx);
		if (rc != EBT_OK) return rc;
		const int rcf = syncFrameTable(ctx);
		if (rcf != EBT_OK) return rcf;
	}
	int wpr = 0;
	{
		const int rc = ensureMask(ctx, n, &wpr);
		if (rc != EBT_OK) return rc;
	}
	estimateLoad(ctx);
	for (int i = 0; i < P; ++i) ctx->h_order[i] = i;
	std::stable_sort(ctx->h_order, ctx->h_order + P,
					 [&](int a, int b) { return ctx->load_est[static_cast<size_t>(a)] > ctx->load_est[static_cast<size_t>(b)]; });
	CU(cudaMemcpyAsync(ctx->d_order, ctx->h_order, static_cast<size_t>(P) * sizeof(int), cudaMemcpyHostToDevice,
					   ctx->stream));
	CU(cudaMemsetAsync(ctx->d_nrec, 0, sizeof(unsigned long long), ctx->stream));
	tic(ctx);
	CU(ebt_launch_track_batch(ctx->d_patches, P, ctx->d_order, ctx->d_rings, ctx->d_nablas, d_events, n,
							  ctx->event_index, ctx->d_mask, wpr, ctx->prefilter_margin, frameTable(ctx),
							  ctx->dp, ctx->d_records, ctx->d_nrec,
							  static_cast<unsigned long long>(ctx->prm.record_capacity), ctx->stream));
	toc(ctx, K_TRACK);
	finishRecords(ctx);  // host side of the previous push's records, while this push's kernel runs
	return collectPush(ctx, n, ctx->stream, n_opt, false);
}

// EBT_HOST_PROFILE=1: where the host time of a multi-stream push goes (printed at exit; measurement only)
namespace
{
struct HostProfile
{
	bool on = getenv("EBT_HOST_PROFILE") != nullptr;
	double t[8] = {};
	long calls = 0;
	~HostProfile()
	{
		if (!on || !calls) return;
		static const char* names[8] = {"prepare (upload, mask, load estimate)", "sort work + descriptors", "launch calls",
									   "previous records -> host list", "wait for the kernel", "collect (state copy, record copy queued)",
									   "", ""};
		std::fprintf(stderr, "[ebt host profile] %ld pushes, ms per push:", calls);
		for (int i = 0; i < 6; ++i) std::fprintf(stderr, "  %s %.3f;", names[i], t[i] / calls * 1e3);
		std::fprintf(stderr, "\n");
	}
} g_host_profile;
inline double nowSeconds()
{
	timespec ts;
	clock_gettime(CLOCK_MONOTONIC, &ts);
	return ts.tv_sec + ts.tv_nsec * 1e-9;
}
#define HP_MARK(k)                                      \
	if (g_host_profile.on)                              \
	{                                                   \
		const double now_ = nowSeconds();               \
		g_host_profile.t[k] += now_ - hp_t;             \
		hp_t = now_;                                    \
	}
}  // namespace

// One piece of S streams on the shared-pixel-warp kernel: one prefilter launch and
// one stream-kernel launch for all of them, on ctxs[0]'s CUDA stream.
static int runMulti(ebt_ctx* const* ctxs, int S, const ebt_event* const* d_events, const int64_t* n, int64_t* n_opt)
{
	ebt_ctx* ctx = ctxs[0];  // owner of the launch scratch and the CUDA stream used
	cudaStream_t on = ctx->stream;
	double hp_t = g_host_profile.on ? nowSeconds() : 0.0;
	if (g_host_profile.on) g_host_profile.calls++;
	size_t n_work = 0;
	for (int i = 0; i < S; ++i) n_work += ctxs[i]->h.size();
	if (static_cast<size_t>(S) > ctx->descs_cap)
	{
		CU(cudaStreamSynchronize(on));
		cudaFree(ctx->d_descs);
		cudaFreeHost(ctx->h_descs);
		ctx->d_descs = nullptr;
		ctx->h_descs = nullptr;
		ctx->descs_cap = 0;
		const size_t cap = static_cast<size_t>(S) + 8;
		CU(cudaMalloc(&ctx->d_descs, cap * sizeof(StreamDesc)));
		CU(cudaHostAlloc(&ctx->h_descs, cap * sizeof(StreamDesc), cudaHostAllocDefault));
		ctx->descs_cap = cap;
	}
	if (n_work > ctx->work_cap)
	{
		CU(cudaStreamSynchronize(on));
		cudaFree(ctx->d_work);
		cudaFreeHost(ctx->h_work);
		ctx->d_work = nullptr;
		ctx->h_work = nullptr;
		ctx->work_cap = 0;
		const size_t cap = n_work + n_work / 4 + 64;
		CU(cudaMalloc(&ctx->d_work, cap * sizeof(WorkItem)));
		CU(cudaHostAlloc(&ctx->h_work, cap * sizeof(WorkItem), cudaHostAllocDefault));
		ctx->work_cap = cap;
	}
	int64_t max_events = 0;
	int max_patches = 0;
	size_t w = 0;
	ctx->work_load.clear();
	for (int i = 0; i < S; ++i)
	{
		ebt_ctx* c = ctxs[i];
		// the kernel below rewrites the device record buffers: it must start behind the previous push's copies
		if (c->pending_records) CU(cudaStreamWaitEvent(on, c->rec_event, 0));
		// work of this context still in flight on its own CUDA stream comes first
		if (c != ctx)
		{
			if (!c->join_event) CU(cudaEventCreateWithFlags(&c->join_event, cudaEventDisableTiming));
			CU(cudaEventRecord(c->join_event, c->stream));
			CU(cudaStreamWaitEvent(on, c->join_event, 0));
		}
		cudaStream_t own = c->stream;
		c->stream = on;
		int rc = uploadPatches(c);
		if (rc == EBT_OK) rc = syncFrameTable(c);
		int wpr = 0;
		if (rc == EBT_OK) rc = ensureMask(c, n[i], &wpr);
		c->stream = own;
		if (rc != EBT_OK)
		{
			if (c != ctx) ctx->err = c->err;
			return rc;
		}
		CU(cudaMemsetAsync(c->d_nrec, 0, sizeof(unsigned long long), on));
		StreamDesc& D = ctx->h_descs[i];
		D.patches = c->d_patches;
		D.rings = c->d_rings;
		D.nablas = c->d_nablas;
		D.events = d_events[i];
		D.n_events = n[i];
		D.event_base = c->event_index;
		D.mask = c->d_mask;
		D.grids = c->d_frame_ptrs;
		D.records = c->d_records;
		D.n_records = c->d_nrec;
		D.record_cap = static_cast<unsigned long long>(c->prm.record_capacity);
		D.margin = c->prefilter_margin;
		D.n_patches = static_cast<int>(c->h.size());
		D.words_per_row = wpr;
		D.latest = c->latest;
		D.pad_ = 0;
		max_events = std::max(max_events, n[i]);
		max_patches = std::max(max_patches, D.n_patches);
		estimateLoad(c);
		for (int f = 0; f < D.n_patches; ++f)
		{
			ctx->h_work[w++] = WorkItem{i, f};
			ctx->work_load.push_back(c->load_est[static_cast<size_t>(f)]);
		}
	}
	HP_MARK(0);
	{
		// heaviest first over all streams (stable: equal loads keep stream / list order)
		std::vector<int>& idx = ctx->work_idx;
		idx.resize(n_work);
		for (size_t k = 0; k < n_work; ++k) idx[k] = static_cast<int>(k);
		std::stable_sort(idx.begin(), idx.end(),
						 [&](int a, int b) { return ctx->work_load[static_cast<size_t>(a)] > ctx->work_load[static_cast<size_t>(b)]; });
		ctx->work_tmp.assign(ctx->h_work, ctx->h_work + n_work);
		for (size_t k = 0; k < n_work; ++k) ctx->h_work[k] = ctx->work_tmp[static_cast<size_t>(idx[k])];
	}
	CU(cudaMemcpyAsync(ctx->d_descs, ctx->h_descs, static_cast<size_t>(S) * sizeof(StreamDesc), cudaMemcpyHostToDevice, on));
	CU(cudaMemcpyAsync(ctx->d_work, ctx->h_work, n_work * sizeof(WorkItem), cudaMemcpyHostToDevice, on));
	HP_MARK(1);
	tic(ctx);
	CU(ebt_launch_track_multi(ctx->d_descs, S, ctx->d_work, static_cast<int>(n_work), max_events, max_patches, ctx->dp,
							  ctx->d_queue, on));
	toc(ctx, K_TRACK);
	for (int i = 0; i < S; ++i)
	{
		ebt_ctx* c = ctxs[i];
		CU(cudaMemcpyAsync(c->nrec_stage, c->d_nrec, sizeof(unsigned long long), cudaMemcpyDeviceToHost, on));
		CU(cudaMemcpyAsync(c->h_stage, c->d_patches, c->h.size() * sizeof(DevPatch), cudaMemcpyDeviceToHost, on));
	}
	HP_MARK(2);
	// host side of the previous push's records, while this push's kernel runs
	for (int i = 0; i < S; ++i) finishRecords(ctxs[i]);
	HP_MARK(3);
	CU(cudaStreamSynchronize(on));
	HP_MARK(4);
	int result = EBT_OK;
	for (int i = 0; i < S; ++i)
	{
		const int rc = collectPush(ctxs[i], n[i], on, n_opt ? n_opt + i : nullptr, true);
		if (rc != EBT_OK)
		{
			if (ctxs[i] != ctx) ctx->err = ctxs[i]->err;
			result = rc;
		}
	}
	HP_MARK(5);
	return result;
}

// Which kernel a push of these contexts takes: the shared-pixel-warp kernel needs the
// getter-only products and the cost map off and enough features to fill its persistent
// CTAs (more than two per SM); ebt_set_stream_kernel(ctx, 0) forces one CTA per feature.
static bool multiEligible(ebt_ctx* const* ctxs, int S)
{
	size_t feats = 0;
	for (int i = 0; i < S; ++i)
	{
		const ebt_ctx* c = ctxs[i];
		if (c->stream_kernel == 0 || c->dp.products || c->dp.cost_maps) return false;
		feats += c->h.size();
	}
	return S > 1 || ctxs[0]->stream_kernel == 2 || feats > 2 * static_cast<size_t>(ctxs[0]->sm_count);
}

// A push is cut into pieces small enough for the prefilter scratch (256 MiB per
// context) and for the record buffer (expected optimisations per event of the
// previous piece, with a factor 4 in hand); the stream kernel is sequential in the
// events, so the cut changes nothing.
static int64_t pieceEvents(const ebt_ctx* ctx)
{
	const int64_t rows = std::max<int64_t>(1, static_cast<int64_t>(ctx->h.size()));
	const int64_t by_mask = std::max<int64_t>(1 << 16, ((int64_t(1) << 26) / rows) * 32);
	const double by_records = static_cast<double>(ctx->prm.record_capacity) / (4.0 * ctx->opt_per_event);
	return std::max<int64_t>(4096, std::min<int64_t>(by_mask, static_cast<int64_t>(std::min(by_records, 4e18))));
}

static int runPiecesMulti(ebt_ctx* const* ctxs, int S, const ebt_event* const* d_events, const int64_t* n, int64_t* n_opt)
{
	std::vector<int64_t> done(static_cast<size_t>(S), 0), m(static_cast<size_t>(S), 0), k(static_cast<size_t>(S), 0);
	std::vector<ebt_ctx*> act;
	std::vector<const ebt_event*> ev;
	std::vector<int64_t> cnt;
	std::vector<int> who;
	if (n_opt)
		for (int i = 0; i < S; ++i) n_opt[i] = 0;
	for (;;)
	{
		act.clear();
		ev.clear();
		cnt.clear();
		who.clear();
		for (int i = 0; i < S; ++i)
		{
			ebt_ctx* c = ctxs[i];
			const int64_t left = n[i] - done[static_cast<size_t>(i)];
			if (left <= 0) continue;
			if (c->h.empty())
			{
				// no patches: the events only advance the stream position
				c->event_index += left;
				done[static_cast<size_t>(i)] = n[i];
				continue;
			}
			act.push_back(c);
			ev.push_back(d_events[i] + done[static_cast<size_t>(i)]);
			cnt.push_back(std::min(left, pieceEvents(c)));
			who.push_back(i);
		}
		if (act.empty()) return EBT_OK;
		const int A = static_cast<int>(act.size());
		std::vector<int64_t> got(static_cast<size_t>(A), 0);
		int rc;
		if (multiEligible(act.data(), A))
			rc = runMulti(act.data(), A, ev.data(), cnt.data(), got.data());
		else
		{
			rc = EBT_OK;
			for (int a = 0; a < A && rc == EBT_OK; ++a) rc = runStream(act[static_cast<size_t>(a)], ev[static_cast<size_t>(a)], cnt[static_cast<size_t>(a)], &got[static_cast<size_t>(a)]);
		}
		for (int a = 0; a < A; ++a)
		{
			done[static_cast<size_t>(who[static_cast<size_t>(a)])] += cnt[static_cast<size_t>(a)];
			if (n_opt) n_opt[who[static_cast<size_t>(a)]] += got[static_cast<size_t>(a)];
		}
		if (rc != EBT_OK)
		{
			if (act[0] != ctxs[0]) ctxs[0]->err = act[0]->err;
			return rc;
		}
	}
}

static int runPieces(ebt_ctx* ctx, const ebt_event* d_events, int64_t n, int64_t* n_opt)
{
	if (n_opt) *n_opt = 0;
	if (n <= 0) return EBT_OK;
	return runPiecesMulti(&ctx, 1, &d_events, &n, n_opt);
}

// Host-buffer upload of one context's batch into its own device buffer, on its own CUDA stream.
static int uploadEvents(ebt_ctx* ctx, const ebt_event* events, int64_t n)
{
	if (n <= 0) return EBT_OK;
	if (static_cast<size_t>(n) > ctx->d_events_cap)
	{
		CU(cudaStreamSynchronize(ctx->stream));
		cudaFree(ctx->d_events);
		ctx->d_events = nullptr;
		ctx->d_events_cap = 0;
		const size_t cap = static_cast<size_t>(n) + static_cast<size_t>(n) / 4 + 1024;
		CU(cudaMalloc(&ctx->d_events, cap * sizeof(ebt_event)));
		ctx->d_events_cap = cap;
	}
	CU(cudaMemcpyAsync(ctx->d_events, events, static_cast<size_t>(n) * sizeof(ebt_event), cudaMemcpyHostToDevice,
					   ctx->stream));
	return EBT_OK;
}

// FeatureDetector::updatePatches for a batch of events (feature_detector.cpp:585-619).
int ebt_push_events(ebt_ctx* ctx, const ebt_event* events, int64_t n, int64_t* n_optimizations)
{
	if (!ctx || n < 0 || (n > 0 && !events)) return fail(ctx, EBT_E_INVALID, "ebt_push_events: bad argument");
	CU(cudaSetDevice(ctx->device));
	{
		const int rc = uploadEvents(ctx, events, n);
		if (rc != EBT_OK) return rc;
	}
	return runPieces(ctx, ctx->d_events, n, n_optimizations);
}

int ebt_push_events_multi(ebt_ctx* const* ctxs, int n_streams, const ebt_event* const* events, const int64_t* n_events,
						  int device_resident, int64_t* n_optimizations)
{
	ebt_ctx* ctx = (ctxs && n_streams > 0) ? ctxs[0] : nullptr;
	if (!ctx || !events || !n_events) return fail(ctx, EBT_E_INVALID, "ebt_push_events_multi: bad argument");
	for (int i = 0; i < n_streams; ++i)
	{
		if (!ctxs[i] || n_events[i] < 0 || (n_events[i] > 0 && !events[i]))
			return fail(ctx, EBT_E_INVALID, "ebt_push_events_multi: bad argument for stream " + std::to_string(i));
		if (ctxs[i]->device != ctx->device)
			return fail(ctx, EBT_E_INVALID, "ebt_push_events_multi: all contexts must live on the same device");
		// the solver constants travel as one kernel parameter for the whole launch
		if (std::memcmp(&ctxs[i]->dp, &ctx->dp, offsetof(DevParams, products)) != 0)
			return fail(ctx, EBT_E_INVALID, "ebt_push_events_multi: the contexts' parameters (image size, solver "
											"constants) must be equal; stream " + std::to_string(i) + " differs");
		for (int j = 0; j < i; ++j)
			if (ctxs[j] == ctxs[i]) return fail(ctx, EBT_E_INVALID, "ebt_push_events_multi: a context is listed twice");
	}
	CU(cudaSetDevice(ctx->device));
	std::vector<const ebt_event*> dev(static_cast<size_t>(n_streams));
	for (int i = 0; i < n_streams; ++i)
	{
		if (device_resident)
			dev[static_cast<size_t>(i)] = events[i];
		else
		{
			const int rc = uploadEvents(ctxs[i], events[i], n_events[i]);
			if (rc != EBT_OK)
			{
				if (ctxs[i] != ctx) ctx->err = ctxs[i]->err;
				return rc;
			}
			dev[static_cast<size_t>(i)] = ctxs[i]->d_events;
		}
	}
	return runPiecesMulti(ctxs, n_streams, dev.data(), n_events, n_optimizations);
}

int ebt_push_events_device(ebt_ctx* ctx, const ebt_event* d_events, int64_t n, int64_t* n_optimizations)
{
	if (!ctx || n < 0 || (n > 0 && !d_events))
		return fail(ctx, EBT_E_INVALID, "ebt_push_events_device: bad argument");
	CU(cudaSetDevice(ctx->device));
	return runPieces(ctx, d_events, n, n_optimizations);
}

// ---- feature-batch sharding of one stream over several contexts (GPUs) ----------
int ebt_set_shard(ebt_ctx* ctx, int index, int count)
{
	if (!ctx || count < 1 || index < 0 || index >= count)
		return fail(ctx, EBT_E_INVALID, "ebt_set_shard: need 0 <= index < count");
	if (!ctx->h.empty() || !ctx->arch.empty())
		return fail(ctx, EBT_E_INVALID, "ebt_set_shard: set the shard before the first image (or after ebt_reset)");
	ctx->shard_index = index;
	ctx->shard_count = count;
	return EBT_OK;
}

size_t ebt_sizeof_patch_state(void) { return sizeof(DevPatch); }

int ebt_get_patch_state(ebt_ctx* ctx, void* out, int cap, int* n_out)
{
	if (!ctx || !n_out || cap < 0 || (cap > 0 && !out)) return fail(ctx, EBT_E_INVALID, "ebt_get_patch_state: bad argument");
	*n_out = static_cast<int>(ctx->h.size());
	const size_t n = std::min(ctx->h.size(), static_cast<size_t>(cap));
	if (n) std::memcpy(out, ctx->h.data(), n * sizeof(DevPatch));
	return EBT_OK;
}

int ebt_merge_patch_state(ebt_ctx* ctx, const void* const* states, int n_shards, int n_patches)
{
	if (!ctx || !states || n_shards != ctx->shard_count)
		return fail(ctx, EBT_E_INVALID, "ebt_merge_patch_state: one state per shard is needed");
	if (n_patches != static_cast<int>(ctx->h.size()))
		return fail(ctx, EBT_E_INVALID, "ebt_merge_patch_state: the shards' patch lists differ in length");
	for (int s = 0; s < n_shards; ++s)
		if (!states[s] && n_patches > 0) return fail(ctx, EBT_E_INVALID, "ebt_merge_patch_state: null state");
	for (size_t i = 0; i < ctx->h.size(); ++i)
	{
		DevPatch& mine = ctx->h[i];
		const int owner = mine.track_id % ctx->shard_count;
		if (owner == ctx->shard_index) continue;
		DevPatch theirs;
		std::memcpy(&theirs, static_cast<const unsigned char*>(states[owner]) + i * sizeof(DevPatch), sizeof(DevPatch));
		if (theirs.track_id != mine.track_id)
			return fail(ctx, EBT_E_INVALID, "ebt_merge_patch_state: the shards' patch lists are out of step at entry " +
											std::to_string(i));
		// the owner's tracking state; where the patch's data lives on THIS device stays ours
		theirs.store_slot = mine.store_slot;
		theirs.frame_slot = mine.frame_slot;
		theirs.foreign = 1;
		mine = theirs;
	}
	ctx->host_dirty = true;
	return EBT_OK;
}

int ebt_get_tracks(ebt_ctx* ctx, ebt_track* out, int cap, int archived, int* n_out)
{
	if (!ctx || !n_out) return fail(ctx, EBT_E_INVALID, "ebt_get_tracks: bad argument");
	const auto& ps = archived ? ctx->arch : ctx->h;
	for (size_t i = 0; i < ps.size() && static_cast<int>(i) < cap && out; ++i) fillTrack(ps[i], out[i]);
	*n_out = static_cast<int>(ps.size());
	return EBT_OK;
}

static const HostTrack* findTrack(ebt_ctx* ctx, int32_t id, const DevPatch** p, int* archIdx)
{
	materialise(ctx);
	for (size_t i = 0; i < ctx->h.size(); ++i)
		if (ctx->h[i].track_id == id)
		{
			*p = &ctx->h[i];
			*archIdx = -1;
			return &ctx->all_tracks[static_cast<size_t>(id)];
		}
	for (size_t i = 0; i < ctx->arch.size(); ++i)
		if (ctx->arch[i].track_id == id)
		{
			*p = &ctx->arch[i];
			*archIdx = static_cast<int>(i);
			return &ctx->all_tracks[static_cast<size_t>(id)];
		}
	return nullptr;
}

int ebt_get_patch(ebt_ctx* ctx, int32_t track_id, int which, double* out625)
{
	if (!ctx || !out625 || which < 0 || which > 2) return fail(ctx, EBT_E_INVALID, "ebt_get_patch: bad argument");
	if (which != 0 && !ctx->dp.products)
		return fail(ctx, EBT_E_INVALID, "ebt_get_patch: this product is not computed; enable it with ebt_set_products");
	const DevPatch* p = nullptr;
	int ai = -1;
	if (!findTrack(ctx, track_id, &p, &ai)) return fail(ctx, EBT_E_INVALID, "ebt_get_patch: unknown track id");
	const size_t off = static_cast<size_t>(p->store_slot) * EBT_PATCH_PIXELS;
	if (which == 2)
	{
		if (ai >= 0)
		{
			if (ctx->arch_pred[static_cast<size_t>(ai)].empty())
				return fail(ctx, EBT_E_INVALID, "ebt_get_patch: track was archived before the products were enabled");
			std::memcpy(out625, ctx->arch_pred[static_cast<size_t>(ai)].data(), EBT_PATCH_PIXELS * sizeof(double));
		}
		else
		{
			CU(cudaSetDevice(ctx->device));
			CU(cudaMemcpy(out625, ctx->d_pred_nablas + off, EBT_PATCH_PIXELS * sizeof(double), cudaMemcpyDeviceToHost));
		}
		return EBT_OK;
	}
	int nab[EBT_PATCH_PIXELS];
	if (ai >= 0)
	{
		const std::vector<int>& src = which == 0 ? ctx->arch_nabla[static_cast<size_t>(ai)] : ctx->arch_mc[static_cast<size_t>(ai)];
		if (src.empty())
			return fail(ctx, EBT_E_INVALID, "ebt_get_patch: track was archived before the products were enabled");
		std::memcpy(nab, src.data(), sizeof(nab));
	}
	else
	{
		CU(cudaSetDevice(ctx->device));
		CU(cudaMemcpy(nab, (which == 0 ? ctx->d_nablas : ctx->d_mc_nablas) + off, sizeof(nab), cudaMemcpyDeviceToHost));
	}
	for (int i = 0; i < EBT_PATCH_PIXELS; ++i) out625[i] = static_cast<double>(nab[i]);
	return EBT_OK;
}

// DetectorParams::drawImages and the per-optimisation Patch::warpImage /
// integrateMotionCompensatedEvents calls (feature_detector.cpp:504-510, 615;
// optimizer.cpp:182): products that only feed getters.
int ebt_set_products(ebt_ctx* ctx, int mode)
{
	if (!ctx || mode < 0 || mode > 2) return fail(ctx, EBT_E_INVALID, "ebt_set_products: mode must be 0, 1 or 2");
	CU(cudaSetDevice(ctx->device));
	if (mode && !ctx->d_mc_nablas)
	{
		const size_t n = static_cast<size_t>(ctx->prm.max_patches) * EBT_PATCH_PIXELS;
		CU(cudaMalloc(&ctx->d_mc_nablas, n * sizeof(int)));
		CU(cudaMalloc(&ctx->d_pred_nablas, n * sizeof(double)));
		CU(cudaMemset(ctx->d_mc_nablas, 0, n * sizeof(int)));
		CU(cudaMemset(ctx->d_pred_nablas, 0, n * sizeof(double)));
	}
	ctx->dp.products = mode;
	ctx->dp.mc_nablas = ctx->d_mc_nablas;
	ctx->dp.pred_nablas = ctx->d_pred_nablas;
	return EBT_OK;
}

// OptimizerParams::drawCostMap (optimizer.h:18): Optimizer::drawCostMap
// (optimizer.cpp:32-60) after every optimisation that follows.
int ebt_set_cost_map(ebt_ctx* ctx, int on)
{
	if (!ctx) return fail(ctx, EBT_E_INVALID, "ebt_set_cost_map: bad argument");
	CU(cudaSetDevice(ctx->device));
	if (on && !ctx->d_cost_maps)
	{
		const size_t n = static_cast<size_t>(ctx->prm.max_patches) * 121;
		CU(cudaMalloc(&ctx->d_cost_maps, n * sizeof(double)));
		CU(cudaMemset(ctx->d_cost_maps, 0, n * sizeof(double)));
	}
	ctx->dp.cost_maps = on ? ctx->d_cost_maps : nullptr;
	return EBT_OK;
}

// Patch::getCostMap (patch.h:56): 11x11 row-major, map(y + 5, x + 5); zeros
// until the patch was optimised with the cost map on (the reference holds
// zeros(1, 1) then, patch.cpp:29).
int ebt_get_cost_map(ebt_ctx* ctx, int32_t track_id, double* out121)
{
	if (!ctx || !out121) return fail(ctx, EBT_E_INVALID, "ebt_get_cost_map: bad argument");
	if (!ctx->d_cost_maps)
		return fail(ctx, EBT_E_INVALID, "ebt_get_cost_map: not computed; enable it with ebt_set_cost_map");
	const DevPatch* p = nullptr;
	int ai = -1;
	if (!findTrack(ctx, track_id, &p, &ai)) return fail(ctx, EBT_E_INVALID, "ebt_get_cost_map: unknown track id");
	if (ai >= 0)
	{
		const std::vector<double>& src = ctx->arch_cost[static_cast<size_t>(ai)];
		if (src.empty())
			std::memset(out121, 0, 121 * sizeof(double));
		else
			std::memcpy(out121, src.data(), 121 * sizeof(double));
		return EBT_OK;
	}
	CU(cudaSetDevice(ctx->device));
	CU(cudaStreamSynchronize(ctx->stream));
	CU(cudaMemcpy(out121, ctx->d_cost_maps + static_cast<size_t>(p->store_slot) * 121, 121 * sizeof(double),
				  cudaMemcpyDeviceToHost));
	return EBT_OK;
}

// Cost maps of the problems of the last ebt_optimize_batch / _device call made
// with the cost map on: n x 121 doubles.
int ebt_get_batch_cost_maps(ebt_ctx* ctx, double* out, int64_t n)
{
	if (!ctx || !out || n < 0) return fail(ctx, EBT_E_INVALID, "ebt_get_batch_cost_maps: bad argument");
	if (static_cast<size_t>(n) > ctx->n_bcost)
		return fail(ctx, EBT_E_INVALID, "ebt_get_batch_cost_maps: the last batched solve produced fewer cost maps");
	CU(cudaSetDevice(ctx->device));
	CU(cudaStreamSynchronize(ctx->stream));
	CU(cudaMemcpy(out, ctx->d_bcost, static_cast<size_t>(n) * 121 * sizeof(double), cudaMemcpyDeviceToHost));
	return EBT_OK;
}

int ebt_get_trajectory(ebt_ctx* ctx, int32_t track_id, double* out_xyt, int cap, int* n_out)
{
	if (!ctx || !n_out) return fail(ctx, EBT_E_INVALID, "ebt_get_trajectory: bad argument");
	const DevPatch* p = nullptr;
	int ai = -1;
	const HostTrack* t = findTrack(ctx, track_id, &p, &ai);
	if (!t) return fail(ctx, EBT_E_INVALID, "ebt_get_trajectory: unknown track id");
	const int n = static_cast<int>(t->traj.size() / 3);
	if (out_xyt) std::memcpy(out_xyt, t->traj.data(), static_cast<size_t>(std::min(n, cap)) * 3 * sizeof(double));
	*n_out = n;
	return EBT_OK;
}

int ebt_get_final_costs(ebt_ctx* ctx, int32_t track_id, double* out, int cap, int* n_out)
{
	if (!ctx || !n_out) return fail(ctx, EBT_E_INVALID, "ebt_get_final_costs: bad argument");
	const DevPatch* p = nullptr;
	int ai = -1;
	const HostTrack* t = findTrack(ctx, track_id, &p, &ai);
	if (!t) return fail(ctx, EBT_E_INVALID, "ebt_get_final_costs: unknown track id");
	const int n = static_cast<int>(t->costs.size());
	if (out) std::memcpy(out, t->costs.data(), static_cast<size_t>(std::min(n, cap)) * sizeof(double));
	*n_out = n;
	return EBT_OK;
}

int ebt_get_records(ebt_ctx* ctx, ebt_opt_record* out, int64_t cap, int64_t* n_out)
{
	if (!ctx || !n_out) return fail(ctx, EBT_E_INVALID, "ebt_get_records: bad argument");
	sortRecords(ctx);
	const int64_t n = static_cast<int64_t>(ctx->records.size());
	if (out) std::memcpy(out, ctx->records.data(), static_cast<size_t>(std::min(n, cap)) * sizeof(ebt_opt_record));
	*n_out = n;
	return EBT_OK;
}

int ebt_clear_records(ebt_ctx* ctx)
{
	if (!ctx) return EBT_E_INVALID;
	materialise(ctx);  // the journal refers to ranges of records
	ctx->records.clear();
	ctx->records_sorted = 0;
	ctx->record_batches.clear();
	return EBT_OK;
}

static int copyGrid(ebt_ctx* ctx, int slot, double* gx, double* gy)
{
	const size_t npx = static_cast<size_t>(ctx->prm.width) * ctx->prm.height;
	std::vector<double2> g(npx);
	CU(cudaSetDevice(ctx->device));
	CU(cudaMemcpy(g.data(), ctx->frames[slot].grid, npx * sizeof(double2), cudaMemcpyDeviceToHost));
	for (size_t i = 0; i < npx; ++i)
	{
		gx[i] = g[i].x;
		gy[i] = g[i].y;
	}
	return EBT_OK;
}

int ebt_get_gradients(ebt_ctx* ctx, double* gx, double* gy)
{
	if (!ctx || !gx || !gy || !ctx->frames[ctx->latest].grid)
		return fail(ctx, EBT_E_INVALID, "ebt_get_gradients: no frame yet");
	return copyGrid(ctx, ctx->latest, gx, gy);
}

// ---- batched entry -------------------------------------------------------
int ebt_frames_add(ebt_ctx* ctx, const uint8_t* gray)
{
	if (!ctx || !gray) return fail(ctx, EBT_E_INVALID, "ebt_frames_add: bad argument");
	CU(cudaSetDevice(ctx->device));
	int slot = allocFrameSlot(ctx);
	if (slot < 0) return fail(ctx, EBT_E_CAPACITY, "ebt_frames_add: no free template frame slot");
	const int rc = buildGrid(ctx, gray, slot);
	if (rc != EBT_OK) return rc;
	ctx->frames[slot].alive = true;
	ctx->frames[slot].users = 1LL << 40;  // pinned: never released by new_frame_end
	ctx->frames[slot].ts = INT64_MIN + slot;
	return slot;
}

int ebt_frames_get(ebt_ctx* ctx, int slot, double* gx, double* gy)
{
	if (!ctx || slot < 0 || slot >= static_cast<int>(ctx->frames.size()) || !ctx->frames[slot].grid || !gx || !gy)
		return fail(ctx, EBT_E_INVALID, "ebt_frames_get: bad argument");
	return copyGrid(ctx, slot, gx, gy);
}

int ebt_optimize_batch_device(ebt_ctx* ctx, const ebt_problem* d_problems, int64_t n_problems,
							  const ebt_event* d_events, ebt_solution* d_solutions, int32_t* d_nabla)
{
	if (!ctx || n_problems < 0 || n_problems > INT32_MAX || (n_problems > 0 && (!d_problems || !d_events || !d_solutions)))
		return fail(ctx, EBT_E_INVALID, "ebt_optimize_batch_device: bad argument");
	CU(cudaSetDevice(ctx->device));
	DevParams dp = ctx->dp;
	ctx->n_bcost = 0;
	if (dp.cost_maps)
	{
		// the batched kernel indexes its cost maps by problem, not by store slot
		const size_t need = static_cast<size_t>(n_problems) * 121;
		if (need > ctx->cap_bcost)
		{
			CU(cudaStreamSynchronize(ctx->stream));
			cudaFree(ctx->d_bcost);
			ctx->d_bcost = nullptr;
			ctx->cap_bcost = 0;
			CU(cudaMalloc(&ctx->d_bcost, need * sizeof(double)));
			ctx->cap_bcost = need;
		}
		CU(cudaMemsetAsync(ctx->d_bcost, 0, need * sizeof(double), ctx->stream));
		dp.cost_maps = ctx->d_bcost;
		ctx->n_bcost = static_cast<size_t>(n_problems);
	}
	{
		const int rcf = syncFrameTable(ctx);
		if (rcf != EBT_OK) return rcf;
	}
	tic(ctx);
	CU(ebt_launch_optimize_batch(d_problems, static_cast<int>(n_problems), d_events, frameTable(ctx),
								 dp, d_solutions, d_nabla, ctx->stream));
	toc(ctx, K_BATCH);
	return EBT_OK;
}

}  // extern "C"
template <typename T>
static cudaError_t ensure(T** p, size_t* cap, size_t need)
{
	if (need <= *cap) return cudaSuccess;
	cudaFree(*p);
	*p = nullptr;
	*cap = 0;
	cudaError_t e = cudaMalloc(p, need * sizeof(T));
	if (e == cudaSuccess) *cap = need;
	return e;
}
extern "C" {

int ebt_optimize_batch(ebt_ctx* ctx, const ebt_problem* problems, int64_t n_problems,
					   const ebt_event* events, int64_t n_events, ebt_solution* solutions,
					   int32_t* nabla)
{
	if (!ctx || n_problems < 0 || n_events < 0 || (n_problems > 0 && (!problems || !events || !solutions)))
		return fail(ctx, EBT_E_INVALID, "ebt_optimize_batch: bad argument");
	if (n_problems == 0) return EBT_OK;
	for (int64_t i = 0; i < n_problems; ++i)
	{
		const ebt_problem& p = problems[i];
		if (p.n_events < 1 || p.event_offset < 0 || static_cast<int64_t>(p.event_offset) + p.n_events > n_events ||
			p.frame_slot < 0 || p.frame_slot >= static_cast<int>(ctx->frames.size()) || !ctx->frames[p.frame_slot].grid)
			return fail(ctx, EBT_E_INVALID, "ebt_optimize_batch: problem " + std::to_string(i) + " is malformed");
	}
	CU(cudaSetDevice(ctx->device));
	const size_t np = static_cast<size_t>(n_problems);
	{
		CU(ensure(&ctx->d_problems, &ctx->cap_problems, np));
		CU(ensure(&ctx->d_solutions, &ctx->cap_solutions, np));
		CU(ensure(&ctx->d_bevents, &ctx->cap_bevents, static_cast<size_t>(n_events)));
		if (nabla) CU(ensure(&ctx->d_bnabla, &ctx->cap_bnabla, np * EBT_PATCH_PIXELS));
	}
	CU(cudaMemcpyAsync(ctx->d_problems, problems, np * sizeof(ebt_problem), cudaMemcpyHostToDevice, ctx->stream));
	CU(cudaMemcpyAsync(ctx->d_bevents, events, static_cast<size_t>(n_events) * sizeof(ebt_event),
					   cudaMemcpyHostToDevice, ctx->stream));
	const int rc = ebt_optimize_batch_device(ctx, ctx->d_problems, n_problems, ctx->d_bevents,
											 ctx->d_solutions, nabla ? ctx->d_bnabla : nullptr);
	if (rc != EBT_OK) return rc;
	CU(cudaMemcpyAsync(solutions, ctx->d_solutions, np * sizeof(ebt_solution), cudaMemcpyDeviceToHost, ctx->stream));
	if (nabla)
		CU(cudaMemcpyAsync(nabla, ctx->d_bnabla, np * EBT_PATCH_PIXELS * sizeof(int), cudaMemcpyDeviceToHost,
						   ctx->stream));
	CU(cudaStreamSynchronize(ctx->stream));
	return EBT_OK;
}

}  // extern "C"

// ---------------------------------------------------------------------------
// Per-frame re-arm on the device: corner detector and pyramidal
// LK, kernels in ebt_frontend.cu.
// ---------------------------------------------------------------------------
namespace
{
int ensureFrontend(ebt_ctx* ctx)
{
	Frontend& fe = ctx->fe;
	if (fe.ready) return EBT_OK;
	const int W = ctx->prm.width, H = ctx->prm.height;
	const size_t npx = static_cast<size_t>(W) * H;
	// buildOpticalFlowPyramid (lkpyramid.cpp): a level is added only while it
	// stays larger than the window
	fe.levels = 1;
	fe.w[0] = W;
	fe.h[0] = H;
	for (int l = 1; l <= fe.prm.lk_levels && l < EBT_LK_MAX_LEVELS; ++l)
	{
		const int w = (fe.w[l - 1] + 1) / 2, h = (fe.h[l - 1] + 1) / 2;
		if (w <= EBT_LK_WINDOW || h <= EBT_LK_WINDOW) break;
		fe.w[l] = w;
		fe.h[l] = h;
		fe.levels = l + 1;
	}
	for (int l = 0; l < fe.levels; ++l)
	{
		const size_t n = static_cast<size_t>(fe.w[l]) * fe.h[l];
		CU(cudaMalloc(&fe.img[0][l], n));
		CU(cudaMalloc(&fe.img[1][l], n));
		CU(cudaMalloc(&fe.deriv[l], n * sizeof(short2)));
	}
	CU(cudaMalloc(&fe.scratch, npx));
	CU(cudaMalloc(&fe.planes, 3 * npx * sizeof(float)));
	CU(cudaMalloc(&fe.resp, npx * sizeof(float)));
	CU(cudaMalloc(&fe.max_key, sizeof(unsigned)));
	CU(cudaMalloc(&fe.n_cand, sizeof(int)));
	// a 3x3 local maximum excludes its 8 neighbours unless they tie: npx/2 bounds
	// any checkerboard of ties, npx/4 + slack covers images that are not adversarial
	fe.cand_cap = static_cast<int>(npx / 4 + 1024);
	CU(cudaMalloc(&fe.cand, static_cast<size_t>(fe.cand_cap) * sizeof(ebt_corner_candidate)));
	CU(cudaHostAlloc(&fe.h_cand, static_cast<size_t>(fe.cand_cap) * sizeof(ebt_corner_candidate),
					 cudaHostAllocDefault));
	CU(cudaHostAlloc(&fe.h_ncand, sizeof(int), cudaHostAllocDefault));
	fe.ready = true;
	return EBT_OK;
}

int ensureLkPoints(ebt_ctx* ctx, int n)
{
	Frontend& fe = ctx->fe;
	if (n <= fe.lk_cap) return EBT_OK;
	cudaFree(fe.d_pts);
	cudaFree(fe.d_status);
	cudaFreeHost(fe.h_pts);
	cudaFreeHost(fe.h_status);
	fe.d_pts = fe.h_pts = nullptr;
	fe.d_status = fe.h_status = nullptr;
	fe.lk_cap = 0;
	const int cap = std::max(n, 1024);
	CU(cudaMalloc(&fe.d_pts, static_cast<size_t>(cap) * 4 * sizeof(float)));
	CU(cudaMalloc(&fe.d_status, static_cast<size_t>(cap)));
	CU(cudaHostAlloc(&fe.h_pts, static_cast<size_t>(cap) * 4 * sizeof(float), cudaHostAllocDefault));
	CU(cudaHostAlloc(&fe.h_status, static_cast<size_t>(cap), cudaHostAllocDefault));
	fe.lk_cap = cap;
	return EBT_OK;
}

// goodFeaturesToTrack after the response map (featureselect.cpp): candidates
// (thresholded 3x3 maxima inside the mask, found on the device) sorted by
// response, equal responses by descending address (greaterThanPtr), then the
// greedy minimum-distance selection on a grid of minDistance cells.
void selectCorners(const ebt_corner_candidate* cand, int n, int W, int H, int max_corners,
				   double min_distance, std::vector<double>& out)
{
	std::vector<int> idx(static_cast<size_t>(n));
	for (int i = 0; i < n; ++i) idx[static_cast<size_t>(i)] = i;
	std::sort(idx.begin(), idx.end(), [&](int a, int b) {
		const ebt_corner_candidate &ca = cand[a], &cb = cand[b];
		if (ca.value != cb.value) return ca.value > cb.value;
		return ca.y * W + ca.x > cb.y * W + cb.x;
	});
	out.clear();
	if (min_distance >= 1)
	{
		const int cell = static_cast<int>(std::lrint(min_distance));  // cvRound
		const int gw = (W + cell - 1) / cell, gh = (H + cell - 1) / cell;
		std::vector<std::vector<int>> grid(static_cast<size_t>(gw) * gh);  // indices into out / 2
		const double md2 = min_distance * min_distance;
		for (int k : idx)
		{
			const int x = cand[k].x, y = cand[k].y;
			const int xc = x / cell, yc = y / cell;
			const int x1 = std::max(0, xc - 1), y1 = std::max(0, yc - 1);
			const int x2 = std::min(gw - 1, xc + 1), y2 = std::min(gh - 1, yc + 1);
			bool good = true;
			for (int yy = y1; yy <= y2 && good; ++yy)
				for (int xx = x1; xx <= x2 && good; ++xx)
					for (int j : grid[static_cast<size_t>(yy) * gw + xx])
					{
						const double dx = x - out[2 * static_cast<size_t>(j)], dy = y - out[2 * static_cast<size_t>(j) + 1];
						if (dx * dx + dy * dy < md2)
						{
							good = false;
							break;
						}
					}
			if (!good) continue;
			grid[static_cast<size_t>(yc) * gw + xc].push_back(static_cast<int>(out.size() / 2));
			out.push_back(x);
			out.push_back(y);
			if (max_corners > 0 && static_cast<int>(out.size() / 2) == max_corners) break;
		}
	}
	else
	{
		for (int k : idx)
		{
			out.push_back(cand[k].x);
			out.push_back(cand[k].y);
			if (max_corners > 0 && static_cast<int>(out.size() / 2) == max_corners) break;
		}
	}
}

// detectFeatures on a device image
int detectOnDevice(ebt_ctx* ctx, const unsigned char* d_img, std::vector<double>& corners)
{
	Frontend& fe = ctx->fe;
	const int W = ctx->prm.width, H = ctx->prm.height;
	tic(ctx);
	CU(ebt_launch_harris(d_img, W, H, 0.04f, fe.prm.quality_level, EBT_PATCH_EXTENT, fe.planes,
						 fe.resp, fe.max_key, fe.cand, fe.cand_cap, fe.n_cand, ctx->stream));
	toc(ctx, K_CORNERS);
	CU(cudaMemcpyAsync(fe.h_ncand, fe.n_cand, sizeof(int), cudaMemcpyDeviceToHost, ctx->stream));
	CU(cudaStreamSynchronize(ctx->stream));
	const int n = *fe.h_ncand;
	if (n > fe.cand_cap)
		return fail(ctx, EBT_E_CAPACITY, "corner detector: more local maxima than the candidate buffer holds");
	if (n > 0)
	{
		CU(cudaMemcpyAsync(fe.h_cand, fe.cand, static_cast<size_t>(n) * sizeof(ebt_corner_candidate),
						   cudaMemcpyDeviceToHost, ctx->stream));
		CU(cudaStreamSynchronize(ctx->stream));
	}
	// FeatureDetector::reset (feature_detector.cpp:34-36)
	const int max_corners = W * H / (EBT_PATCH_SIDE * EBT_PATCH_SIDE);
	selectCorners(fe.h_cand, n, W, H, max_corners, fe.prm.min_distance, corners);
	return EBT_OK;
}

int buildPyramid(ebt_ctx* ctx, int which)
{
	Frontend& fe = ctx->fe;
	if (fe.built[which] == fe.levels) return EBT_OK;
	tic(ctx);
	for (int l = 1; l < fe.levels; ++l)
		CU(ebt_launch_pyr_down(fe.img[which][l - 1], fe.w[l - 1], fe.h[l - 1], fe.img[which][l], ctx->stream));
	toc(ctx, K_PYRAMID);
	fe.built[which] = fe.levels;
	return EBT_OK;
}

// calcOpticalFlowPyrLK(img[prev], img[cur], h_pts[0..n)) -> h_pts[2n..4n), h_status
int lkOnDevice(ebt_ctx* ctx, int prev, int cur, int n)
{
	Frontend& fe = ctx->fe;
	int rc = buildPyramid(ctx, prev);
	if (rc != EBT_OK) return rc;
	rc = buildPyramid(ctx, cur);
	if (rc != EBT_OK) return rc;
	ebt_lk_pyramids pyr;
	pyr.levels = fe.levels;
	tic(ctx);
	for (int l = 0; l < fe.levels; ++l)
	{
		pyr.w[l] = fe.w[l];
		pyr.h[l] = fe.h[l];
		pyr.prev[l] = fe.img[prev][l];
		pyr.cur[l] = fe.img[cur][l];
		pyr.deriv[l] = fe.deriv[l];
		CU(ebt_launch_scharr(fe.img[prev][l], fe.w[l], fe.h[l], fe.deriv[l], ctx->stream));
	}
	toc(ctx, K_PYRAMID);
	const size_t nb = static_cast<size_t>(n) * 2 * sizeof(float);
	CU(cudaMemcpyAsync(fe.d_pts, fe.h_pts, nb, cudaMemcpyHostToDevice, ctx->stream));
	tic(ctx);
	CU(ebt_launch_lk(pyr, fe.d_pts, n, fe.d_pts + 2 * static_cast<size_t>(n), fe.d_status, fe.prm.lk_max_count,
					 fe.prm.lk_epsilon, ctx->stream));
	toc(ctx, K_LK);
	CU(cudaMemcpyAsync(fe.h_pts + 2 * static_cast<size_t>(n), fe.d_pts + 2 * static_cast<size_t>(n), nb,
					   cudaMemcpyDeviceToHost, ctx->stream));
	CU(cudaMemcpyAsync(fe.h_status, fe.d_status, static_cast<size_t>(n), cudaMemcpyDeviceToHost, ctx->stream));
	CU(cudaStreamSynchronize(ctx->stream));
	return EBT_OK;
}
}  // namespace

extern "C" {

int ebt_default_frontend_params(ebt_frontend_params* p)
{
	if (!p) return EBT_E_INVALID;
	std::memset(p, 0, sizeof(*p));
	p->quality_level = 0.01;
	p->min_distance = 10.0;
	p->lk_levels = 3;
	p->lk_max_count = 30;
	p->lk_epsilon = 0.01;
	return EBT_OK;
}

int ebt_set_frontend_params(ebt_ctx* ctx, const ebt_frontend_params* p)
{
	if (!ctx || !p || !(p->quality_level > 0) || p->min_distance < 0 || p->lk_levels < 0 ||
		p->lk_levels >= EBT_LK_MAX_LEVELS || p->lk_max_count < 1 || !(p->lk_epsilon >= 0))
		return fail(ctx, EBT_E_INVALID, "ebt_set_frontend_params: bad argument");
	if (ctx->fe.ready && p->lk_levels != ctx->fe.prm.lk_levels)
		return fail(ctx, EBT_E_INVALID, "ebt_set_frontend_params: lk_levels is fixed once a frame has been seen");
	ctx->fe.prm = *p;
	return EBT_OK;
}

int ebt_detect_features(ebt_ctx* ctx, const uint8_t* gray, double* cornersXY, int cap, int* n_out)
{
	if (!ctx || !gray || !n_out || cap < 0 || (cap > 0 && !cornersXY))
		return fail(ctx, EBT_E_INVALID, "ebt_detect_features: bad argument");
	CU(cudaSetDevice(ctx->device));
	int rc = ensureFrontend(ctx);
	if (rc != EBT_OK) return rc;
	Frontend& fe = ctx->fe;
	const size_t npx = static_cast<size_t>(ctx->prm.width) * ctx->prm.height;
	CU(cudaMemcpyAsync(fe.scratch, gray, npx, cudaMemcpyHostToDevice, ctx->stream));
	std::vector<double> corners;
	rc = detectOnDevice(ctx, fe.scratch, corners);
	if (rc != EBT_OK) return rc;
	const int n = static_cast<int>(corners.size() / 2);
	*n_out = n;
	if (cap > 0) std::memcpy(cornersXY, corners.data(), static_cast<size_t>(std::min(n, cap)) * 2 * sizeof(double));
	return EBT_OK;
}

int ebt_get_features(ebt_ctx* ctx, double* cornersXY, int cap, int* n_out)
{
	if (!ctx || !n_out || cap < 0 || (cap > 0 && !cornersXY))
		return fail(ctx, EBT_E_INVALID, "ebt_get_features: bad argument");
	const int n = static_cast<int>(ctx->corners.size() / 2);
	*n_out = n;
	if (cap > 0 && n > 0)
		std::memcpy(cornersXY, ctx->corners.data(), static_cast<size_t>(std::min(n, cap)) * 2 * sizeof(double));
	return EBT_OK;
}

int ebt_lk_flow(ebt_ctx* ctx, const uint8_t* prev_gray, const uint8_t* cur_gray, const float* ptsXY, int n,
				float* nextXY, uint8_t* status)
{
	if (!ctx || !prev_gray || !cur_gray || n < 0 || (n > 0 && (!ptsXY || !nextXY || !status)))
		return fail(ctx, EBT_E_INVALID, "ebt_lk_flow: bad argument");
	if (n == 0) return EBT_OK;
	CU(cudaSetDevice(ctx->device));
	int rc = ensureFrontend(ctx);
	if (rc != EBT_OK) return rc;
	Frontend& fe = ctx->fe;
	if (fe.have[0] || fe.have[1])
		return fail(ctx, EBT_E_INVALID, "ebt_lk_flow: not available between ebt_new_image calls; use a second context");
	rc = ensureLkPoints(ctx, n);
	if (rc != EBT_OK) return rc;
	const size_t npx = static_cast<size_t>(ctx->prm.width) * ctx->prm.height;
	CU(cudaMemcpyAsync(fe.img[0][0], prev_gray, npx, cudaMemcpyHostToDevice, ctx->stream));
	CU(cudaMemcpyAsync(fe.img[1][0], cur_gray, npx, cudaMemcpyHostToDevice, ctx->stream));
	fe.built[0] = fe.built[1] = 0;
	std::memcpy(fe.h_pts, ptsXY, static_cast<size_t>(n) * 2 * sizeof(float));
	rc = lkOnDevice(ctx, 0, 1, n);
	if (rc != EBT_OK) return rc;
	std::memcpy(nextXY, fe.h_pts + 2 * static_cast<size_t>(n), static_cast<size_t>(n) * 2 * sizeof(float));
	std::memcpy(status, fe.h_status, static_cast<size_t>(n));
	return EBT_OK;
}

int ebt_new_image(ebt_ctx* ctx, const uint8_t* gray, int64_t ts_us)
{
	if (!ctx || !gray) return fail(ctx, EBT_E_INVALID, "ebt_new_image: bad argument");
	CU(cudaSetDevice(ctx->device));
	int rc = ensureFrontend(ctx);
	if (rc != EBT_OK) return rc;
	Frontend& fe = ctx->fe;
	const size_t npx = static_cast<size_t>(ctx->prm.width) * ctx->prm.height;
	// FlowEstimator::addImage (flow_estimator.cpp:16-25): current becomes previous
	fe.cur ^= 1;
	const int cur = fe.cur, prev = cur ^ 1;
	CU(cudaMemcpyAsync(fe.img[cur][0], gray, npx, cudaMemcpyHostToDevice, ctx->stream));
	fe.built[cur] = 0;
	fe.have[cur] = true;
	// extractPatches: detectFeatures, gradients, associatePatches
	std::vector<double> corners;
	rc = detectOnDevice(ctx, fe.img[cur][0], corners);  // synchronises: `gray` may be reused by the caller
	if (rc != EBT_OK) return rc;
	rc = newFrameBegin(ctx, fe.img[cur][0], ts_us, corners.data(), static_cast<int>(corners.size() / 2), true);
	if (rc != EBT_OK) return rc;
	// getFlowPatches (flow_estimator.cpp:27-82)
	if (ctx->image_counter == 2 && fe.have[prev])
	{
		int n = 0;
		for (const DevPatch& p : ctx->h) n += p.init ? 0 : 1;
		if (n > 0)
		{
			rc = ensureLkPoints(ctx, n);
			if (rc != EBT_OK) return rc;
			int k = 0;
			for (const DevPatch& p : ctx->h)
			{
				if (p.init) continue;
				double cx, cy;
				cornerOf(p, cx, cy);
				fe.h_pts[2 * k] = static_cast<float>(cx);  // getFlow(const cv::Point2f&)
				fe.h_pts[2 * k + 1] = static_cast<float>(cy);
				++k;
			}
			rc = lkOnDevice(ctx, prev, cur, n);
			if (rc != EBT_OK) return rc;
			k = 0;
			for (DevPatch& p : ctx->h)
			{
				if (p.init) continue;
				applyFlow(ctx, p, fe.h_status[k], fe.h_pts[2 * (n + k)], fe.h_pts[2 * (n + k) + 1]);
				++k;
			}
		}
	}
	return ebt_new_frame_end(ctx);
}

}  // extern "C"

// ---------------------------------------------------------------------------
// Tracker -> VO hand-off
// ---------------------------------------------------------------------------
static int landScratch(ebt_ctx* ctx, size_t n)
{
	Frontend& fe = ctx->fe;
	if (n <= fe.land_cap) return EBT_OK;
	cudaFree(fe.d_land);
	fe.d_land = nullptr;
	fe.land_cap = 0;
	const size_t cap = std::max<size_t>(n, 1024);
	CU(cudaMalloc(&fe.d_land, cap * 8 * sizeof(double)));  // 2 in + 2 corner + 3 bearing + 1 (id)
	fe.land_cap = cap;
	return EBT_OK;
}

extern "C" {

int ebt_get_landmarks(ebt_ctx* ctx, const ebt_camera* cam, int32_t* track_ids, double* cornersXY,
					  double* bearingsXYZ, int cap, int* n_out)
{
	if (!ctx || !n_out || cap < 0 || (cap > 0 && (!track_ids || !cornersXY)) || (bearingsXYZ && !cam))
		return fail(ctx, EBT_E_INVALID, "ebt_get_landmarks: bad argument");
	const int n = static_cast<int>(ctx->h.size());
	*n_out = n;
	const int m = std::min(n, cap);
	if (m == 0) return EBT_OK;
	CU(cudaSetDevice(ctx->device));
	int rc = uploadPatches(ctx);  // the device copy is the source
	if (rc != EBT_OK) return rc;
	rc = landScratch(ctx, static_cast<size_t>(m));
	if (rc != EBT_OK) return rc;
	double* d = ctx->fe.d_land;
	const size_t M = static_cast<size_t>(m);
	double* d_corner = d + 2 * ctx->fe.land_cap;
	double* d_bear = d + 4 * ctx->fe.land_cap;
	int* d_ids = reinterpret_cast<int*>(d + 7 * ctx->fe.land_cap);
	ebt_camera c{};
	if (cam) c = *cam;
	CU(ebt_launch_unproject(ctx->d_patches, nullptr, m, c, bearingsXYZ ? 1 : 0, d_ids, d_corner, d_bear, ctx->stream));
	CU(cudaMemcpyAsync(track_ids, d_ids, M * sizeof(int), cudaMemcpyDeviceToHost, ctx->stream));
	CU(cudaMemcpyAsync(cornersXY, d_corner, M * 2 * sizeof(double), cudaMemcpyDeviceToHost, ctx->stream));
	if (bearingsXYZ) CU(cudaMemcpyAsync(bearingsXYZ, d_bear, M * 3 * sizeof(double), cudaMemcpyDeviceToHost, ctx->stream));
	CU(cudaStreamSynchronize(ctx->stream));
	return EBT_OK;
}

int ebt_unproject(ebt_ctx* ctx, const ebt_camera* cam, const double* pointsXY, int n, double* bearingsXYZ)
{
	if (!ctx || !cam || n < 0 || (n > 0 && (!pointsXY || !bearingsXYZ)))
		return fail(ctx, EBT_E_INVALID, "ebt_unproject: bad argument");
	if (n == 0) return EBT_OK;
	CU(cudaSetDevice(ctx->device));
	const int rc = landScratch(ctx, static_cast<size_t>(n));
	if (rc != EBT_OK) return rc;
	double* d = ctx->fe.d_land;
	double* d_bear = d + 4 * ctx->fe.land_cap;
	const size_t N = static_cast<size_t>(n);
	CU(cudaMemcpyAsync(d, pointsXY, N * 2 * sizeof(double), cudaMemcpyHostToDevice, ctx->stream));
	CU(ebt_launch_unproject(nullptr, d, n, *cam, 1, nullptr, nullptr, d_bear, ctx->stream));
	CU(cudaMemcpyAsync(bearingsXYZ, d_bear, N * 3 * sizeof(double), cudaMemcpyDeviceToHost, ctx->stream));
	CU(cudaStreamSynchronize(ctx->stream));
	return EBT_OK;
}

}  // extern "C"

// ---------------------------------------------------------------------------
// Event ingestion: text parser on the device, Replayer order
// ---------------------------------------------------------------------------
extern "C" {

int ebt_parse_events_text(ebt_ctx* ctx, const char* text, size_t len, int final_chunk, ebt_event* out,
						  int64_t cap, int64_t* n_out, size_t* consumed)
{
	if (!ctx || !n_out || cap < 0 || (len > 0 && !text) || (cap > 0 && !out))
		return fail(ctx, EBT_E_INVALID, "ebt_parse_events_text: bad argument");
	*n_out = 0;
	if (consumed) *consumed = 0;
	if (len == 0) return EBT_OK;
	if (len > (size_t(1) << 33))
		return fail(ctx, EBT_E_INVALID, "ebt_parse_events_text: pass the text in chunks of at most 8 GiB");
	CU(cudaSetDevice(ctx->device));
	const size_t n_tiles = static_cast<size_t>(ebt_ingest_tiles(len));
	if (len + 16 > ctx->text_cap)
	{
		cudaFree(ctx->d_text);
		ctx->d_text = nullptr;
		ctx->text_cap = 0;
		CU(cudaMalloc(&ctx->d_text, len + len / 4 + 4096));
		ctx->text_cap = len + len / 4 + 4096;
	}
	if (n_tiles > ctx->tiles_cap)
	{
		cudaFree(ctx->d_tile_counts);
		ctx->d_tile_counts = nullptr;
		ctx->tiles_cap = 0;
		CU(cudaMalloc(&ctx->d_tile_counts, (n_tiles + n_tiles / 4 + 64) * sizeof(int)));
		ctx->tiles_cap = n_tiles + n_tiles / 4 + 64;
	}
	if (!ctx->d_total)
	{
		CU(cudaMalloc(&ctx->d_total, 2 * sizeof(long long)));
		CU(cudaHostAlloc(&ctx->h_total, 2 * sizeof(long long), cudaHostAllocDefault));
	}
	CU(cudaMemcpyAsync(ctx->d_text, text, len, cudaMemcpyHostToDevice, ctx->stream));
	CU(cudaMemsetAsync(ctx->d_total + 1, 0xff, sizeof(long long), ctx->stream));
	tic(ctx);
	CU(ebt_launch_count_lines(ctx->d_text, len, ctx->d_tile_counts, ctx->d_total, ctx->stream));
	CU(cudaMemcpyAsync(ctx->h_total, ctx->d_total, sizeof(long long), cudaMemcpyDeviceToHost, ctx->stream));
	CU(cudaStreamSynchronize(ctx->stream));
	const long long newlines = ctx->h_total[0];
	// a final chunk's last line needs no newline (std::getline); a non-final chunk's
	// unterminated tail is left to the caller (*consumed)
	const bool tail = final_chunk && text[len - 1] != '\n';
	const long long n_lines = newlines + (tail ? 1 : 0);
	*n_out = n_lines;
	if (n_lines > cap)
	{
		toc(ctx, K_PARSE);
		return fail(ctx, EBT_E_CAPACITY, "ebt_parse_events_text: " + std::to_string(n_lines) + " lines, output holds " +
											 std::to_string(cap));
	}
	if (n_lines > 0)
	{
		if (static_cast<size_t>(n_lines) > ctx->parsed_cap)
		{
			cudaFree(ctx->d_parsed);
			ctx->d_parsed = nullptr;
			ctx->parsed_cap = 0;
			const size_t c = static_cast<size_t>(n_lines) + static_cast<size_t>(n_lines) / 4 + 1024;
			CU(cudaMalloc(&ctx->d_parsed, c * sizeof(ebt_event)));
			ctx->parsed_cap = c;
		}
		CU(ebt_launch_parse_lines(ctx->d_text, len, ctx->d_tile_counts, n_lines, final_chunk ? 1 : 0, ctx->d_parsed,
								  reinterpret_cast<unsigned long long*>(ctx->d_total + 1), ctx->stream));
		toc(ctx, K_PARSE);
		CU(cudaMemcpyAsync(ctx->h_total + 1, ctx->d_total + 1, sizeof(long long), cudaMemcpyDeviceToHost, ctx->stream));
		CU(cudaMemcpyAsync(out, ctx->d_parsed, static_cast<size_t>(n_lines) * sizeof(ebt_event), cudaMemcpyDeviceToHost,
						   ctx->stream));
		CU(cudaStreamSynchronize(ctx->stream));
		const unsigned long long e = static_cast<unsigned long long>(ctx->h_total[1]);
		if (e != ~0ull)
		{
			const unsigned code = static_cast<unsigned>(e & 0xff);
			const char* why = code == 2   ? "timestamp has too many digits for an exact conversion (mantissa >= 2^53)"
							  : code == 3 ? "line longer than 128 bytes"
										  : "malformed line (expected \"t_sec x y p\" with p 0 or 1)";
			*n_out = 0;
			return fail(ctx, EBT_E_INVALID, "ebt_parse_events_text: line " + std::to_string(e >> 8) + ": " + why);
		}
	}
	else
		toc(ctx, K_PARSE);
	if (consumed)
	{
		if (final_chunk)
			*consumed = len;
		else
		{
			size_t k = len;
			while (k > 0 && text[k - 1] != '\n') --k;
			*consumed = k;
		}
	}
	return EBT_OK;
}

// tools::Replayer::next (tools/replayer/src/replayer.cpp:56-98) over in-memory
// samples: an event is delivered while its timestamp is < the next image's,
// otherwise the image; replay ends when either runs out (finished(), :40-43).
//
// The whole recording is known here, so the event upload is pipelined (SURVEY 8f n2,
// "chunked async H2D"): while the stream kernel of batch k runs on the context's
// stream, batch k + 1 travels host -> pinned staging -> device in 4 MiB chunks on a
// second CUDA stream (ping-pong device buffers, one CUDA event per batch orders the
// kernel behind its own upload).  Results equal ebt_push_events + ebt_new_image.
namespace
{
constexpr int64_t kReplayChunk = int64_t(1) << 18;  // events per H2D chunk (4 MiB)

int replayBuffers(ebt_ctx* ctx, int64_t max_batch)
{
	if (!ctx->copy_stream)
	{
		CU(cudaStreamCreateWithFlags(&ctx->copy_stream, cudaStreamNonBlocking));
		for (int i = 0; i < 2; ++i) CU(cudaEventCreateWithFlags(&ctx->uploaded[i], cudaEventDisableTiming));
	}
	if (static_cast<size_t>(max_batch) > ctx->replay_cap)
	{
		CU(cudaStreamSynchronize(ctx->copy_stream));
		CU(cudaStreamSynchronize(ctx->stream));
		const size_t cap = static_cast<size_t>(max_batch) + static_cast<size_t>(max_batch) / 8 + 1024;
		for (int i = 0; i < 2; ++i)
		{
			cudaFree(ctx->d_replay[i]);
			ctx->d_replay[i] = nullptr;
			cudaFreeHost(ctx->h_replay[i]);
			ctx->h_replay[i] = nullptr;
		}
		ctx->replay_cap = 0;
		for (int i = 0; i < 2; ++i)
		{
			CU(cudaMalloc(&ctx->d_replay[i], cap * sizeof(ebt_event)));
			CU(cudaHostAlloc(&ctx->h_replay[i], static_cast<size_t>(2 * kReplayChunk) * sizeof(ebt_event), cudaHostAllocDefault));
		}
		ctx->replay_cap = cap;
	}
	return EBT_OK;
}

// host events [lo, hi) -> device buffer `slot`, chunk by chunk on the copy stream; the
// pinned staging of a slot holds two chunks, so the host copy of a chunk overlaps the
// DMA of the one before it
int stageBatch(ebt_ctx* ctx, int slot, const ebt_event* events, int64_t lo, int64_t hi)
{
	int half = 0;
	for (int64_t at = lo; at < hi; at += kReplayChunk, half ^= 1)
	{
		const int64_t m = std::min(kReplayChunk, hi - at);
		ebt_event* stage = ctx->h_replay[slot] + static_cast<size_t>(half) * kReplayChunk;
		// this half's previous DMA (two chunks ago) must have left the staging buffer
		if (at - lo >= 2 * kReplayChunk) CU(cudaEventSynchronize(ctx->chunk_done[slot][half]));
		std::memcpy(stage, events + at, static_cast<size_t>(m) * sizeof(ebt_event));
		CU(cudaMemcpyAsync(ctx->d_replay[slot] + (at - lo), stage, static_cast<size_t>(m) * sizeof(ebt_event),
						   cudaMemcpyHostToDevice, ctx->copy_stream));
		if (!ctx->chunk_done[slot][half]) CU(cudaEventCreateWithFlags(&ctx->chunk_done[slot][half], cudaEventDisableTiming));
		CU(cudaEventRecord(ctx->chunk_done[slot][half], ctx->copy_stream));
	}
	CU(cudaEventRecord(ctx->uploaded[slot], ctx->copy_stream));
	return EBT_OK;
}
}  // namespace

int ebt_replay(ebt_ctx* ctx, const ebt_event* events, int64_t n_events, const uint8_t* frames,
			   const int64_t* frame_ts, int n_frames, int64_t* n_events_delivered, int64_t* n_optimizations)
{
	if (!ctx || n_events < 0 || n_frames < 0 || (n_events > 0 && !events) || (n_frames > 0 && (!frames || !frame_ts)))
		return fail(ctx, EBT_E_INVALID, "ebt_replay: bad argument");
	if (n_events_delivered) *n_events_delivered = 0;
	if (n_optimizations) *n_optimizations = 0;
	CU(cudaSetDevice(ctx->device));
	const size_t npx = static_cast<size_t>(ctx->prm.width) * ctx->prm.height;
	// the batch in front of every image: events strictly older than it
	std::vector<int64_t> cut(static_cast<size_t>(n_frames) + 1, 0);
	int64_t max_batch = 0;
	for (int k = 0; k < n_frames; ++k)
	{
		int64_t lo = cut[static_cast<size_t>(k)], hi = n_events;
		while (lo < hi)
		{
			const int64_t mid = lo + (hi - lo) / 2;
			if (events[mid].t_us < frame_ts[k])
				lo = mid + 1;
			else
				hi = mid;
		}
		cut[static_cast<size_t>(k) + 1] = lo;
		max_batch = std::max(max_batch, lo - cut[static_cast<size_t>(k)]);
	}
	if (max_batch > 0)
	{
		const int rc = replayBuffers(ctx, max_batch);
		if (rc != EBT_OK) return rc;
	}
	int64_t pos = 0, n_opt = 0;
	int staged = -1;  // batch (image index) already on its way to the device
	auto stage = [&](int k) -> int {
		if (k >= n_frames || cut[static_cast<size_t>(k) + 1] <= cut[static_cast<size_t>(k)]) return EBT_OK;
		staged = k;
		return stageBatch(ctx, k & 1, events, cut[static_cast<size_t>(k)], cut[static_cast<size_t>(k) + 1]);
	};
	for (int k = 0; k < n_frames && pos < n_events; ++k)
	{
		const int64_t lo = cut[static_cast<size_t>(k)], hi = cut[static_cast<size_t>(k) + 1];
		if (hi > lo)
		{
			if (staged != k)
			{
				const int rc = stage(k);
				if (rc != EBT_OK) return rc;
			}
			CU(cudaStreamWaitEvent(ctx->stream, ctx->uploaded[k & 1], 0));
			// the next batch travels while this one is tracked
			int next = k + 1;
			while (next < n_frames && cut[static_cast<size_t>(next) + 1] <= cut[static_cast<size_t>(next)]) ++next;
			if (next < n_frames && (next & 1) != (k & 1))
			{
				const int rc = stage(next);
				if (rc != EBT_OK) return rc;
			}
			int64_t m = 0;
			const int rc = runPieces(ctx, ctx->d_replay[k & 1], hi - lo, &m);
			n_opt += m;
			if (rc != EBT_OK) return rc;
			pos = hi;
		}
		if (pos >= n_events) break;  // hasEvents_ is false before the image is reached
		const int rc = ebt_new_image(ctx, frames + static_cast<size_t>(k) * npx, frame_ts[k]);
		if (rc != EBT_OK) return rc;
	}
	if (ctx->copy_stream) CU(cudaStreamSynchronize(ctx->copy_stream));
	if (n_events_delivered) *n_events_delivered = pos;
	if (n_optimizations) *n_optimizations = n_opt;
	return EBT_OK;
}

}  // extern "C"
