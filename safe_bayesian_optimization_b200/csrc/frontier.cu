// frontier.cu -- the two O(m) host loops that sit right after the safe mask in the reference, on the GPU.
//
// (1) sbo_frontier_image: the binary image FindSafetyContourIndices hands to cv::findContours
//     (reference src/safe_bayesian_optimization_node.cpp:431-456) with the reference's own pixel mapping --
//     bounds truncated to int (:431-434), pixel = int((coord - min) / (max - min) * dim) (:450-451), cells whose
//     pixel falls outside dropped (:453), a later cell overwriting an earlier one on the same pixel -- plus the
//     pixel -> cell lookup tables that replace the nested unordered_map of all m cells (:468-475, the slowest
//     O(m) piece of the node).  On a regular grid both maps are separable (one table per axis), so the tables are
//     W + H integers built on the host and the image is a gather  img[py][px] = mask[row_of_py[py]][col_of_px[px]].
//     cv::findContours itself (Suzuki-Abe border following, inherently sequential) stays on the host.
// (2) sbo_compact_cells: index list of the safe cells in ascending cell order (replaces the gather loop in front of
//     the alpha-shape stage, :944-949) or only of the safe cells that touch an unsafe cell / the grid frame in
//     their 8-neighbourhood (a superset of every pixel border following can visit; interior cells cannot be
//     alpha-shape boundary vertices once alpha >= the grid pitch), so the CGAL stage sees O(sqrt m) points.
//     Deterministic order: per-block counts, one-block exclusive scan, ordered scatter.
//
// Both kernels are HBM-bound byte/integer work: 1 B read + 1 B written per pixel for the image, 1 B (safe list) or
// ~1 B through L1/L2-served neighbour reads (border list) per cell plus 4 B per emitted index.
#include <algorithm>

#include "common.cuh"

namespace sbo {
namespace {

constexpr int kFrThreads = 256;
constexpr int kFrCellsPerBlock = 4096;  // 16 cells per thread

__global__ void __launch_bounds__(kFrThreads)
frontier_image_kernel(const uint8_t* __restrict__ mask, int W, int H, const int32_t* __restrict__ col_of_px,
                      const int32_t* __restrict__ row_of_py, uint8_t* __restrict__ img) {
  const int py = blockIdx.y;
  const int src_row = row_of_py[py];
  for (int px = blockIdx.x * kFrThreads + threadIdx.x; px < W; px += gridDim.x * kFrThreads) {
    uint8_t v = 0;
    if (src_row >= 0) {
      const int src_col = col_of_px[px];
      if (src_col >= 0) v = mask[(size_t)src_row * W + src_col] ? 255 : 0;
    }
    img[(size_t)py * W + px] = v;
  }
}

// border = safe and (on the frame of the global grid, or an unsafe cell among the 8 neighbours).  Rows outside this
// shard that are inside the grid are unknown: treated as unsafe (superset).
__device__ __forceinline__ bool cell_selected(const uint8_t* __restrict__ mask, int W, int rows, int row0, int H,
                                              long long c, int mode) {
  if (!mask[c]) return false;
  if (mode == 0) return true;
  const int r = (int)(c / W), x = (int)(c - (long long)r * W);
  if (x == 0 || x == W - 1 || row0 + r == 0 || row0 + r == H - 1) return true;
  if (r == 0 || r == rows - 1) return true;  // shard edge inside the grid
  const uint8_t* p = mask + c;
  const bool all = p[-1] && p[1] && p[-W - 1] && p[-W] && p[-W + 1] && p[W - 1] && p[W] && p[W + 1];
  return !all;
}

__global__ void __launch_bounds__(kFrThreads)
compact_count_kernel(const uint8_t* __restrict__ mask, int W, int rows, int row0, int H, long long m, int mode,
                     unsigned int* __restrict__ block_counts) {
  __shared__ unsigned int warp_sum[kFrThreads / 32];
  const long long base = (long long)blockIdx.x * kFrCellsPerBlock;
  unsigned int cnt = 0;
#pragma unroll 4
  for (int i = 0; i < kFrCellsPerBlock / kFrThreads; ++i) {
    const long long c = base + i * kFrThreads + threadIdx.x;
    if (c < m && cell_selected(mask, W, rows, row0, H, c, mode)) ++cnt;
  }
#pragma unroll
  for (int d = 16; d > 0; d >>= 1) cnt += __shfl_down_sync(0xffffffffu, cnt, d);
  if ((threadIdx.x & 31) == 0) warp_sum[threadIdx.x >> 5] = cnt;
  __syncthreads();
  if (threadIdx.x == 0) {
    unsigned int t = 0;
    for (int w = 0; w < kFrThreads / 32; ++w) t += warp_sum[w];
    block_counts[blockIdx.x] = t;
  }
}

// exclusive scan of the block counts in place (one CTA walks them in chunks of 1024); total -> *total_out
__global__ void __launch_bounds__(1024)
compact_scan_kernel(unsigned int* __restrict__ counts, int nblocks, unsigned long long* __restrict__ total_out) {
  __shared__ unsigned int warp_tot[32];
  __shared__ unsigned long long carry_s;
  if (threadIdx.x == 0) carry_s = 0;
  __syncthreads();
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  for (int b0 = 0; b0 < nblocks; b0 += 1024) {
    const int i = b0 + threadIdx.x;
    const unsigned int v = i < nblocks ? counts[i] : 0u;
    unsigned int incl = v;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
      const unsigned int t = __shfl_up_sync(0xffffffffu, incl, d);
      if (lane >= d) incl += t;
    }
    if (lane == 31) warp_tot[warp] = incl;
    __syncthreads();
    if (warp == 0) {
      unsigned int w = warp_tot[lane], wi = w;
#pragma unroll
      for (int d = 1; d < 32; d <<= 1) {
        const unsigned int t = __shfl_up_sync(0xffffffffu, wi, d);
        if (lane >= d) wi += t;
      }
      warp_tot[lane] = wi - w;  // exclusive prefix of the warp totals
    }
    __syncthreads();
    const unsigned long long carry = carry_s;
    const unsigned long long excl = carry + warp_tot[warp] + (incl - v);
    if (i < nblocks) counts[i] = (unsigned int)excl;  // < 2^31 cells per shard
    __syncthreads();
    if (threadIdx.x == 1023) carry_s = excl + v;
    __syncthreads();
  }
  if (threadIdx.x == 0) *total_out = carry_s;
}

__global__ void __launch_bounds__(kFrThreads)
compact_write_kernel(const uint8_t* __restrict__ mask, int W, int rows, int row0, int H, long long m, int mode,
                     long long cell0, const unsigned int* __restrict__ block_offsets, int32_t* __restrict__ out,
                     long long cap) {
  __shared__ unsigned int warp_base[kFrThreads / 32];
  __shared__ unsigned int running;
  if (threadIdx.x == 0) running = block_offsets[blockIdx.x];
  __syncthreads();
  const long long base = (long long)blockIdx.x * kFrCellsPerBlock;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  for (int i = 0; i < kFrCellsPerBlock / kFrThreads; ++i) {  // ascending cell order: row i of the block's 16 x 256 cells
    const long long c = base + i * kFrThreads + threadIdx.x;
    const bool sel = c < m && cell_selected(mask, W, rows, row0, H, c, mode);
    const unsigned int ballot = __ballot_sync(0xffffffffu, sel);
    if (lane == 0) warp_base[warp] = __popc(ballot);
    __syncthreads();
    unsigned int before = 0, total = 0;
#pragma unroll
    for (int w = 0; w < kFrThreads / 32; ++w) {
      const unsigned int t = warp_base[w];
      if (w < warp) before += t;
      total += t;
    }
    const unsigned int pos = running + before + __popc(ballot & ((1u << lane) - 1u));
    if (sel && (long long)pos < cap) out[pos] = (int32_t)(cell0 + c);
    __syncthreads();
    if (threadIdx.x == 0) running += total;
    __syncthreads();
  }
}

}  // namespace

int32_t frontier_image(sbo_ctx* ctx, const int32_t* col_of_px, const int32_t* row_of_py, uint8_t* img_host) {
  const int W = ctx->W, H = ctx->H;
  if (ctx->row0 != 0 || ctx->row1 != H)
    return fail(ctx, SBO_ERR_STATE, "sbo_frontier_image: the context must own every grid row (the pixel map is global)");
  const size_t m = (size_t)W * H;
  if (ctx->fr_cap < (long long)m) {
    if (ctx->d_fr_img) cudaFree(ctx->d_fr_img);
    if (ctx->d_fr_lut) cudaFree(ctx->d_fr_lut);
    ctx->d_fr_img = nullptr; ctx->d_fr_lut = nullptr; ctx->fr_cap = 0;
    SBO_CUDA(ctx, cudaMalloc(&ctx->d_fr_img, m));
    SBO_CUDA(ctx, cudaMalloc(&ctx->d_fr_lut, sizeof(int32_t) * ((size_t)W + H)));
    ctx->fr_cap = (long long)m;
  }
  SBO_CUDA(ctx, cudaMemcpyAsync(ctx->d_fr_lut, col_of_px, sizeof(int32_t) * W, cudaMemcpyHostToDevice, ctx->stream));
  SBO_CUDA(ctx, cudaMemcpyAsync(ctx->d_fr_lut + W, row_of_py, sizeof(int32_t) * H, cudaMemcpyHostToDevice, ctx->stream));
  dim3 grid((unsigned)std::min((W + kFrThreads - 1) / kFrThreads, 64), (unsigned)H);
  frontier_image_kernel<<<grid, kFrThreads, 0, ctx->stream>>>(ctx->d_mask, W, H, ctx->d_fr_lut, ctx->d_fr_lut + W, ctx->d_fr_img);
  SBO_LAUNCH_CHECK(ctx, "frontier_image_kernel");
  SBO_CUDA(ctx, cudaMemcpyAsync(img_host, ctx->d_fr_img, m, cudaMemcpyDeviceToHost, ctx->stream));
  SBO_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
  return SBO_OK;
}

int32_t compact_cells(sbo_ctx* ctx, int mode, int32_t* idx_host, int64_t cap, int64_t* count_out) {
  const long long m = ctx->m_local;
  *count_out = 0;
  if (m == 0) return SBO_OK;
  if (ctx->cell0 + m > 0x7fffffffLL) return fail(ctx, SBO_ERR_CAPACITY, "sbo_compact_cells: cell indices exceed int32");
  const int nblocks = (int)((m + kFrCellsPerBlock - 1) / kFrCellsPerBlock);
  if (ctx->cp_cap < m) {
    if (ctx->d_cp_idx) cudaFree(ctx->d_cp_idx);
    if (ctx->d_cp_counts) cudaFree(ctx->d_cp_counts);
    ctx->d_cp_idx = nullptr; ctx->d_cp_counts = nullptr; ctx->cp_cap = 0;
    SBO_CUDA(ctx, cudaMalloc(&ctx->d_cp_idx, sizeof(int32_t) * (size_t)m));
    SBO_CUDA(ctx, cudaMalloc(&ctx->d_cp_counts, sizeof(unsigned int) * (size_t)nblocks + sizeof(unsigned long long) * 2));
    ctx->cp_cap = m;
  }
  // the 8-byte total sits behind the counts, 8-byte aligned
  unsigned long long* d_total =
      reinterpret_cast<unsigned long long*>(reinterpret_cast<uintptr_t>(ctx->d_cp_counts + nblocks + 1) & ~(uintptr_t)7);
  const int rows = ctx->row1 - ctx->row0;
  compact_count_kernel<<<nblocks, kFrThreads, 0, ctx->stream>>>(ctx->d_mask, ctx->W, rows, ctx->row0, ctx->H, m, mode,
                                                               ctx->d_cp_counts);
  SBO_LAUNCH_CHECK(ctx, "compact_count_kernel");
  compact_scan_kernel<<<1, 1024, 0, ctx->stream>>>(ctx->d_cp_counts, nblocks, d_total);
  SBO_LAUNCH_CHECK(ctx, "compact_scan_kernel");
  compact_write_kernel<<<nblocks, kFrThreads, 0, ctx->stream>>>(ctx->d_mask, ctx->W, rows, ctx->row0, ctx->H, m, mode, ctx->cell0,
                                                               ctx->d_cp_counts, ctx->d_cp_idx, m);
  SBO_LAUNCH_CHECK(ctx, "compact_write_kernel");
  unsigned long long total = 0;
  SBO_CUDA(ctx, cudaMemcpyAsync(&total, d_total, sizeof(total), cudaMemcpyDeviceToHost, ctx->stream));
  SBO_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
  *count_out = (int64_t)total;
  const int64_t ncopy = std::min<int64_t>((int64_t)total, cap);
  if (idx_host && ncopy > 0) {
    SBO_CUDA(ctx, cudaMemcpyAsync(idx_host, ctx->d_cp_idx, sizeof(int32_t) * (size_t)ncopy, cudaMemcpyDeviceToHost, ctx->stream));
    SBO_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
  }
  return SBO_OK;
}

}  // namespace sbo
