// ebt_kernels.h — launchers of the sm_100a kernels in ebt_kernels.cu.
#pragma once
#include <cuda_runtime.h>

#include "ebt_device_types.h"

// Prefilter + stream kernel.  `order`: feature index per CTA (heaviest first);
// `mask`: n_patches x words_per_row scratch words (ebt_mask_words_per_row);
// `margin`: how far (px) a rect may drift from its position at launch before
// its CTA rebuilds its candidate row.
cudaError_t ebt_launch_track_batch(DevPatch* patches, int n_patches, const int* order,
								   ebt_event* rings, int* nablas, const ebt_event* events,
								   long long n_events, long long event_base, unsigned* mask,
								   int words_per_row, double margin, const FrameTable& ft,
								   const DevParams& P, ebt_opt_record* records,
								   unsigned long long* n_records, unsigned long long record_cap,
								   cudaStream_t stream);
// The same for the event batches of n_streams streams in one launch (shared-pixel-warp
// kernel, persistent CTAs pulling (stream, feature) work items from `queue`).
cudaError_t ebt_launch_track_multi(const StreamDesc* descs, int n_streams, const WorkItem* work, int n_work,
								   long long max_events, int max_patches, const DevParams& P, int* queue,
								   cudaStream_t stream);
int ebt_mask_words_per_row(long long n_events);
cudaError_t ebt_launch_optimize_batch(const ebt_problem* problems, int n_problems,
									  const ebt_event* events, const FrameTable& ft,
									  const DevParams& P, ebt_solution* solutions, int* nablas,
									  cudaStream_t stream);
cudaError_t ebt_launch_num_events(DevPatch* patches, int n_patches, const FrameTable& ft,
								  const DevParams& P, cudaStream_t stream);
cudaError_t ebt_launch_grad(const unsigned char* gray, const double* lut, int W, int H,
							double2* grid, cudaStream_t stream);
cudaError_t ebt_launch_fma_peak(int fp64, void* out, int blocks, int threads, int iters,
								cudaStream_t stream);
// leader-warp phase clocks (all zero unless built with -DEBT_PHASE_CLOCKS); resets them
cudaError_t ebt_launch_phase_clocks(unsigned long long* out8, cudaStream_t stream);
