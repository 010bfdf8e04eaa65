// ebt_ingest.cu — event ingestion on the device: the text
// event format of the reference's dataset reader
//   "t_sec x y p\n"   (/root/reference/tools/dataset_reader/src/davis240c_reader.cpp:60-92)
// parsed into 16-byte ebt_event records, replacing the per-line std::string /
// std::stod / std::stoi parser.
//
// HBM-bound byte work.  Three launches over 4 KiB tiles of the text:
//   newline_count_kernel  one uint4 (16 B) per thread, per-tile '\n' count
//   tile_scan_kernel      exclusive scan of the tile counts (one CTA)
//   parse_lines_kernel    tile staged in shared memory (+ look-ahead), every
//                         thread ranks the newlines in its 16 bytes and parses
//                         the lines that start behind them; line k -> out[k]
// Exactness: the timestamp is std::stod (correctly rounded) followed by
// duration_cast<microseconds>(duration<double>) = (int64)(sec * 1e6).  With the
// decimal mantissa M < 2^53 and 10^f exact (f <= 22), (double)M / 10^f IS the
// correctly rounded value, so the result equals the reference's bit for bit
// (including 0.000249 -> 248); longer mantissas are reported, never guessed.
#include <cuda_runtime.h>
#include <stdint.h>

#include "../../include/ebt_types.h"
#include "ebt_ingest.h"

namespace
{
constexpr int TILE = 4096;
constexpr int THREADS = 256;
constexpr int LOOKAHEAD = 128;  // longest line that may straddle a tile end

__device__ __forceinline__ unsigned newline_bits(uint4 v)
{
	// bit k set when byte k of the 16 is '\n'
	unsigned m = 0;
	const unsigned w[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
	for (int i = 0; i < 4; ++i)
	{
		const unsigned eq = __vcmpeq4(w[i], 0x0a0a0a0au);  // 0xff per equal byte
		m |= ((eq & 1u) | ((eq >> 7) & 2u) | ((eq >> 14) & 4u) | ((eq >> 21) & 8u)) << (4 * i);
	}
	return m;
}

__device__ __forceinline__ uint4 load16(const char* text, size_t len, size_t off)
{
	// text is 16-byte aligned (cudaMalloc); the tail is padded with zeros
	if (off + 16 <= len) return *reinterpret_cast<const uint4*>(text + off);
	unsigned char b[16];
#pragma unroll
	for (int i = 0; i < 16; ++i) b[i] = off + i < len ? (unsigned char)text[off + i] : 0;
	return *reinterpret_cast<uint4*>(b);
}

__global__ void __launch_bounds__(THREADS)
newline_count_kernel(const char* __restrict__ text, size_t len, int n_tiles, int* __restrict__ counts)
{
	__shared__ int wsum[THREADS / 32];
	for (int tile = blockIdx.x; tile < n_tiles; tile += gridDim.x)
	{
		const size_t off = (size_t)tile * TILE + (size_t)threadIdx.x * 16;
		int c = off < len ? __popc(newline_bits(load16(text, len, off))) : 0;
#pragma unroll
		for (int o = 16; o > 0; o >>= 1) c += __shfl_xor_sync(0xffffffffu, c, o);
		if ((threadIdx.x & 31) == 0) wsum[threadIdx.x >> 5] = c;
		__syncthreads();
		if (threadIdx.x == 0)
		{
			int t = 0;
#pragma unroll
			for (int w = 0; w < THREADS / 32; ++w) t += wsum[w];
			counts[tile] = t;
		}
		__syncthreads();
	}
}

// exclusive scan of counts[0..n) in place, total to *total; one CTA of 1024 threads
__global__ void __launch_bounds__(1024)
tile_scan_kernel(int* __restrict__ counts, int n, long long* __restrict__ total)
{
	__shared__ long long wsum[32];
	__shared__ long long carry;
	if (threadIdx.x == 0) carry = 0;
	__syncthreads();
	const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
	for (int base = 0; base < n; base += 1024)
	{
		const int i = base + threadIdx.x;
		const long long v = i < n ? counts[i] : 0;
		long long incl = v;
#pragma unroll
		for (int o = 1; o < 32; o <<= 1)
		{
			const long long up = __shfl_up_sync(0xffffffffu, incl, o);
			if (lane >= o) incl += up;
		}
		if (lane == 31) wsum[warp] = incl;
		__syncthreads();
		long long before = carry;
		for (int w = 0; w < warp; ++w) before += wsum[w];
		if (i < n) counts[i] = (int)(before + incl - v);  // line indices of one call fit 31 bits (checked by the host)
		__syncthreads();
		if (threadIdx.x == 1023) carry = before + incl;
		__syncthreads();
	}
	if (threadIdx.x == 0) *total = carry;
}

__device__ __forceinline__ bool is_digit(int c) { return c >= '0' && c <= '9'; }

// One line "t_sec x y p" starting at s[0]; `avail` bytes are readable.  Returns
// 0 ok, 1 malformed, 2 mantissa too long for an exact conversion, 3 line does
// not end within `avail` (and the buffer goes on).
__device__ int parse_line(const unsigned char* s, int avail, bool buffer_ends_here, ebt_event& e)
{
	int p = 0;
	unsigned long long M = 0;
	int nd = 0, nf = 0;
	while (p < avail && is_digit(s[p]))
	{
		if (nd < 19) M = M * 10 + (s[p] - '0');
		++nd;
		++p;
	}
	if (p < avail && s[p] == '.')
	{
		++p;
		while (p < avail && is_digit(s[p]))
		{
			if (nd < 19) M = M * 10 + (s[p] - '0');
			++nd;
			++nf;
			++p;
		}
	}
	if (nd == 0) return p >= avail && !buffer_ends_here ? 3 : 1;
	if (nd > 19 || M >= (1ull << 53) || nf > 22) return 2;
	double p10 = 1.0;
	for (int i = 0; i < nf; ++i) p10 *= 10.0;  // exact up to 1e22
	const double sec = __ddiv_rn((double)M, p10);
	e.t_us = (long long)__dmul_rn(sec, 1e6);  // duration_cast: truncation
	int v[2];
#pragma unroll
	for (int k = 0; k < 2; ++k)
	{
		if (!(p < avail && s[p] == ' ')) return p >= avail && !buffer_ends_here ? 3 : 1;
		++p;
		int n = 0, x = 0;
		while (p < avail && is_digit(s[p]))
		{
			if (n < 9) x = x * 10 + (s[p] - '0');
			++n;
			++p;
		}
		if (n == 0 || n > 5 || x > 65535) return p >= avail && !buffer_ends_here ? 3 : 1;
		v[k] = x;
	}
	if (!(p < avail && s[p] == ' ')) return p >= avail && !buffer_ends_here ? 3 : 1;
	++p;
	if (p >= avail) return buffer_ends_here ? 1 : 3;
	const int sign = s[p];
	if (sign != '0' && sign != '1') return 1;  // "Sign is not equal to 0/1" (davis240c_reader.cpp:85-88)
	++p;
	if (p < avail && s[p] == '\r') ++p;
	if (p < avail ? s[p] != '\n' : !buffer_ends_here) return p >= avail ? 3 : 1;
	e.x = (uint16_t)v[0];
	e.y = (uint16_t)v[1];
	e.polarity = sign == '1' ? 1 : -1;
	e.pad_[0] = e.pad_[1] = e.pad_[2] = 0;
	return 0;
}

// err[0] = first bad line (min), err[1] = its code
__global__ void __launch_bounds__(THREADS)
parse_lines_kernel(const char* __restrict__ text, size_t len, int n_tiles, const int* __restrict__ tile_first,
				   long long n_lines, int final_chunk, ebt_event* __restrict__ out,
				   unsigned long long* __restrict__ err)
{
	__shared__ __align__(16) unsigned char buf[TILE + LOOKAHEAD];
	__shared__ int wsum[THREADS / 32];
	const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
	for (int tile = blockIdx.x; tile < n_tiles; tile += gridDim.x)
	{
		const size_t base = (size_t)tile * TILE;
		const size_t off = base + (size_t)threadIdx.x * 16;
		const uint4 v = off < len ? load16(text, len, off) : make_uint4(0, 0, 0, 0);
		*reinterpret_cast<uint4*>(&buf[threadIdx.x * 16]) = v;
		if (threadIdx.x < LOOKAHEAD / 16)
		{
			const size_t o2 = base + TILE + (size_t)threadIdx.x * 16;
			*reinterpret_cast<uint4*>(&buf[TILE + threadIdx.x * 16]) =
				o2 < len ? load16(text, len, o2) : make_uint4(0, 0, 0, 0);
		}
		unsigned bits = off < len ? newline_bits(v) : 0u;  // zero padding holds no newline
		const int c = __popc(bits);
		int incl = c;
#pragma unroll
		for (int o = 1; o < 32; o <<= 1)
		{
			const int up = __shfl_up_sync(0xffffffffu, incl, o);
			if (lane >= o) incl += up;
		}
		if (lane == 31) wsum[warp] = incl;
		__syncthreads();
		int before = incl - c;
		for (int w = 0; w < warp; ++w) before += wsum[w];
		// line index of the line that starts right after the k-th newline of this tile
		long long line = (long long)tile_first[tile] + before + 1;
		const int valid = (int)min((size_t)(TILE + LOOKAHEAD), len - base);
		const bool ends = base + valid >= len && final_chunk;
		// the very first line of the buffer starts at offset 0
		if (tile == 0 && threadIdx.x == 0 && n_lines > 0)
		{
			ebt_event e;
			const int rc = parse_line(buf, valid, ends, e);
			if (rc == 0)
				out[0] = e;
			else
				atomicMin(err, (unsigned long long)rc);
		}
		while (bits)
		{
			const int b = __ffs(bits) - 1;
			bits &= bits - 1;
			const int start = threadIdx.x * 16 + b + 1;  // local offset of the line behind this newline
			const long long idx = line++;
			if (idx >= n_lines) continue;  // unterminated tail of a non-final chunk, or nothing behind the last newline
			ebt_event e;
			const int rc = parse_line(buf + start, valid - start, ends, e);
			if (rc == 0)
				out[idx] = e;
			else
				atomicMin(err, ((unsigned long long)idx << 8) | (unsigned)rc);
		}
		__syncthreads();
	}
}
}  // namespace

int ebt_ingest_tiles(size_t len) { return (int)((len + TILE - 1) / TILE); }

cudaError_t ebt_launch_count_lines(const char* d_text, size_t len, int* d_tile_counts, long long* d_total,
								   cudaStream_t stream)
{
	const int n_tiles = ebt_ingest_tiles(len);
	if (n_tiles == 0) return cudaSuccess;
	const int grid = n_tiles < 148 * 8 ? n_tiles : 148 * 8;
	newline_count_kernel<<<grid, THREADS, 0, stream>>>(d_text, len, n_tiles, d_tile_counts);
	tile_scan_kernel<<<1, 1024, 0, stream>>>(d_tile_counts, n_tiles, d_total);
	return cudaGetLastError();
}

cudaError_t ebt_launch_parse_lines(const char* d_text, size_t len, const int* d_tile_first, long long n_lines,
								   int final_chunk, ebt_event* d_out, unsigned long long* d_err,
								   cudaStream_t stream)
{
	const int n_tiles = ebt_ingest_tiles(len);
	if (n_tiles == 0 || n_lines == 0) return cudaSuccess;
	const int grid = n_tiles < 148 * 8 ? n_tiles : 148 * 8;
	parse_lines_kernel<<<grid, THREADS, 0, stream>>>(d_text, len, n_tiles, d_tile_first, n_lines, final_chunk,
													 d_out, d_err);
	return cudaGetLastError();
}

// ---------------------------------------------------------------------------
// ebt_parse_events_text_static — the same parser for callers that must not
// allocate (CUDA graph capture set-up, tight serving loops, harnesses that forbid
// cudaMalloc).  No counterpart in the reference.
// The tile counts, the line total, the error word and the
// tail description live in a file-scope __device__ array of fixed size, so the
// call issues kernels, two small memsets and two small D2H copies and nothing
// else: no cudaMalloc / cudaFree / cudaHostAlloc, no library routine that
// allocates.  Text and output are device buffers the caller owns.
//
// The array is per device (every GPU that loads the library holds its 65 792
// bytes until exit) and shared by every call: it keeps the previous call's
// contents, so the words that must start from a known value (error word, tail
// words) are re-initialised on the stream before use, and the tile counts of
// this call's tiles are all rewritten by newline_count_kernel.  Calls that
// overlap on the GPU (two host threads, or two streams) race on it.
#include <mutex>
#include <stdio.h>

#include "../../include/ebt.h"

namespace
{
constexpr size_t STATIC_MAX_BYTES = EBT_STATIC_PARSE_MAX_BYTES;
constexpr size_t STATIC_MAX_TILES = STATIC_MAX_BYTES / TILE;
constexpr size_t STATIC_HEADER_BYTES = 256;  // [0] newline total, [1] error word, [2] end of last newline, [3] last byte
constexpr int STATIC_MAX_DEVICES = 64;

static __device__ __align__(256) unsigned char g_static_scratch[STATIC_HEADER_BYTES + STATIC_MAX_TILES * sizeof(int)];

std::mutex g_static_mu;
unsigned char* g_static_ptr[STATIC_MAX_DEVICES];  // device address of g_static_scratch, per device

// In host code `g_static_scratch` is only a handle: the device address has to be
// asked for, once per device, with that device current.
cudaError_t static_scratch(int device, unsigned char** out)
{
	std::lock_guard<std::mutex> lock(g_static_mu);
	if (!g_static_ptr[device])
	{
		void* p = nullptr;
		const cudaError_t rc = cudaGetSymbolAddress(&p, g_static_scratch);
		if (rc != cudaSuccess) return rc;
		g_static_ptr[device] = static_cast<unsigned char*>(p);
	}
	*out = g_static_ptr[device];
	return cudaSuccess;
}

// info[2] = offset just behind the last '\n' of the buffer (0 if none), info[3] = last byte
__global__ void tail_info_kernel(const char* __restrict__ text, size_t len, unsigned long long* __restrict__ info)
{
	const int lane = threadIdx.x;
	if (lane == 0) info[3] = (unsigned char)text[len - 1];
	size_t end = len;
	while (end > 0)
	{
		const bool inside = (size_t)lane < end;  // lane k looks at byte end - 1 - k
		const bool nl = inside && text[end - 1 - lane] == '\n';
		const unsigned m = __ballot_sync(0xffffffffu, nl);
		if (m)
		{
			if (lane == 0) info[2] = end - (__ffs(m) - 1);
			return;
		}
		end = end > 32 ? end - 32 : 0;
	}
	if (lane == 0) info[2] = 0;
}

int static_fail(char* errmsg, size_t cap, int code, const char* fmt, long long a = 0, long long b = 0)
{
	if (errmsg && cap) snprintf(errmsg, cap, fmt, a, b);
	return code;
}
}  // namespace

extern "C" int ebt_parse_events_text_static(int device, void* cuda_stream, const char* d_text, size_t len,
											int final_chunk, ebt_event* d_out, int64_t cap, int64_t* n_out,
											size_t* consumed, char* errmsg, size_t errmsg_cap)
{
	if (errmsg && errmsg_cap) errmsg[0] = 0;
	if (!n_out || cap < 0 || (len > 0 && !d_text) || (cap > 0 && !d_out) || device < 0 || device >= STATIC_MAX_DEVICES)
		return static_fail(errmsg, errmsg_cap, EBT_E_INVALID, "ebt_parse_events_text_static: bad argument");
	*n_out = 0;
	if (consumed) *consumed = 0;
	if (len == 0) return EBT_OK;
	if (len > STATIC_MAX_BYTES)
		return static_fail(errmsg, errmsg_cap, EBT_E_CAPACITY,
						   "ebt_parse_events_text_static: %lld bytes of text, the static scratch serves at most %lld "
						   "per call (EBT_STATIC_PARSE_MAX_BYTES); pass the text in smaller chunks",
						   (long long)len, (long long)STATIC_MAX_BYTES);
	if (reinterpret_cast<uintptr_t>(d_text) & 15)
		return static_fail(errmsg, errmsg_cap, EBT_E_INVALID, "ebt_parse_events_text_static: d_text must be 16-byte aligned");
	cudaStream_t stream = static_cast<cudaStream_t>(cuda_stream);
	int prev = -1;
	cudaError_t rc = cudaGetDevice(&prev);
	if (rc == cudaSuccess && prev != device) rc = cudaSetDevice(device);
	unsigned char* scratch = nullptr;
	if (rc == cudaSuccess) rc = static_scratch(device, &scratch);
	unsigned long long host[4] = {0, 0, 0, 0};
	long long n_lines = 0;
	int status = EBT_OK;
	if (rc == cudaSuccess)
	{
		unsigned long long* info = reinterpret_cast<unsigned long long*>(scratch);
		int* counts = reinterpret_cast<int*>(scratch + STATIC_HEADER_BYTES);
		// the array holds the previous call's words: error word to ~0, totals and tail words to 0
		rc = cudaMemsetAsync(info, 0, STATIC_HEADER_BYTES, stream);
		if (rc == cudaSuccess) rc = cudaMemsetAsync(info + 1, 0xff, sizeof(unsigned long long), stream);
		if (rc == cudaSuccess)
			rc = ebt_launch_count_lines(d_text, len, counts, reinterpret_cast<long long*>(info), stream);
		if (rc == cudaSuccess)
		{
			tail_info_kernel<<<1, 32, 0, stream>>>(d_text, len, info);
			rc = cudaGetLastError();
		}
		if (rc == cudaSuccess) rc = cudaMemcpyAsync(host, info, sizeof host, cudaMemcpyDeviceToHost, stream);
		if (rc == cudaSuccess) rc = cudaStreamSynchronize(stream);
		if (rc == cudaSuccess)
		{
			const bool tail = final_chunk && host[3] != '\n';
			n_lines = (long long)host[0] + (tail ? 1 : 0);
			*n_out = n_lines;
			if (n_lines > cap)
				status = static_fail(errmsg, errmsg_cap, EBT_E_CAPACITY,
									 "ebt_parse_events_text_static: %lld lines, output holds %lld", n_lines, (long long)cap);
			else if (n_lines > 0)
			{
				rc = ebt_launch_parse_lines(d_text, len, counts, n_lines, final_chunk ? 1 : 0, d_out, info + 1, stream);
				if (rc == cudaSuccess)
					rc = cudaMemcpyAsync(host + 1, info + 1, sizeof(unsigned long long), cudaMemcpyDeviceToHost, stream);
				if (rc == cudaSuccess) rc = cudaStreamSynchronize(stream);
				if (rc == cudaSuccess && host[1] != ~0ull)
				{
					const unsigned code = (unsigned)(host[1] & 0xff);
					*n_out = 0;
					status = static_fail(errmsg, errmsg_cap, EBT_E_INVALID,
										 code == 2   ? "ebt_parse_events_text_static: line %lld: timestamp has too many digits "
													   "for an exact conversion (mantissa >= 2^53)"
										 : code == 3 ? "ebt_parse_events_text_static: line %lld: line longer than 128 bytes"
													 : "ebt_parse_events_text_static: line %lld: malformed line (expected "
													   "\"t_sec x y p\" with p 0 or 1)",
										 (long long)(host[1] >> 8));
				}
			}
		}
	}
	if (rc != cudaSuccess)
	{
		if (errmsg && errmsg_cap)
			snprintf(errmsg, errmsg_cap, "ebt_parse_events_text_static: CUDA error: %s", cudaGetErrorString(rc));
		status = EBT_E_CUDA;
	}
	else if (status == EBT_OK && consumed)
		*consumed = final_chunk ? len : (size_t)host[2];
	if (prev >= 0 && prev != device) cudaSetDevice(prev);
	return status;
}
