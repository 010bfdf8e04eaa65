// ebt_ingest.h — launchers of the event-text parser kernels in ebt_ingest.cu.
#pragma once
#include <cuda_runtime.h>
#include <stddef.h>

struct ebt_event;

int ebt_ingest_tiles(size_t len);
// per-tile newline counts -> exclusive scan in place (d_tile_counts[t] = newlines before tile t),
// *d_total = newlines in the buffer
cudaError_t ebt_launch_count_lines(const char* d_text, size_t len, int* d_tile_counts, long long* d_total,
								   cudaStream_t stream);
// parses lines 0..n_lines); d_err must hold ~0ull on entry, receives min((line << 8) | code)
cudaError_t ebt_launch_parse_lines(const char* d_text, size_t len, const int* d_tile_first, long long n_lines,
								   int final_chunk, ebt_event* d_out, unsigned long long* d_err,
								   cudaStream_t stream);
