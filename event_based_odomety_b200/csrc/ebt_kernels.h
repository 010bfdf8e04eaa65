// ebt_kernels.h — launchers of the sm_100a kernels in ebt_kernels.cu.
#pragma once
#include <cuda_runtime.h>

#include "ebt_device_types.h"

cudaError_t ebt_launch_track_batch(DevPatch* patches, int n_patches, ebt_event* rings, int* nablas,
								   const ebt_event* events, long long n_events,
								   long long event_base, const FrameTable& ft, const DevParams& P,
								   ebt_opt_record* records, unsigned long long* n_records,
								   unsigned long long record_cap, cudaStream_t stream);
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
