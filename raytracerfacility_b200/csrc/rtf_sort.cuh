// rtf_sort.cuh — stable LSD radix sort of (key, u32 value) pairs, hand-written for sm_100a.
//
// "Onesweep" organisation (Adinets & Merrill 2022): ONE histogram kernel reads the keys once and counts the digits
// of every pass; each pass is then ONE kernel that ranks a tile of 3072 keys, obtains the tile's global offset per
// digit by decoupled look-back over the preceding tiles' digit counts, reorders the tile in shared memory and writes
// out digit runs (coalesced).  A pass moves 12 B in + 12 B out per 63-bit pair: the sort is an HBM-bound streaming
// kernel; 8-bit digits -> 8 passes for 63-bit Morton codes, 4 for 30-bit codes.
//
// Used by the LBVH builder (Morton code, reference id) and by the wavefront ray sort (octant|cell key, queue slot).
// Integer-only and stable, so the result is bit-identical to the CPU oracle's std::stable_sort.
#pragma once
#include "rtf_internal.h"

namespace rtf {

#ifndef RTF_SORT_MATCH
#define RTF_SORT_MATCH 0  // 1: rank with match.any instead of eight ballots
#endif
constexpr int kSortBlock = 256;
constexpr int kSortWarps = kSortBlock / 32;
#ifndef RTF_SORT_ITEMS
#define RTF_SORT_ITEMS 12  // keys per thread: 3072-key tiles (46 KB of shared memory) -> 4 blocks per SM; 16 / 3 blocks: 0.60 ms, 12 / 4: 0.51, 8 / 5: 0.53 (C2 scene, 8 passes)
#endif
constexpr int kSortItems = RTF_SORT_ITEMS;
#ifndef RTF_SORT_MIN_BLOCKS
#define RTF_SORT_MIN_BLOCKS 4
#endif
constexpr int kSortTile = kSortBlock * kSortItems;  // 3072 keys per tile
constexpr int kSortMaxPasses = 8;
constexpr uint32_t kSortFlagAgg = 1u << 30, kSortFlagPrefix = 2u << 30, kSortValueMask = (1u << 30) - 1u;

// ---- 1. digit histograms of all passes in one read ---------------------------------------------------------------
template <class K>
__global__ void __launch_bounds__(256) k_sort_hist(const K* __restrict__ keys, uint32_t n, int passes, int first_bit,
                                                   uint32_t* __restrict__ hist /* passes x 256 */) {
  __shared__ uint32_t sh[kSortMaxPasses * 256];
  for (int i = threadIdx.x; i < passes * 256; i += blockDim.x) sh[i] = 0u;
  __syncthreads();
  for (uint32_t idx = blockIdx.x * blockDim.x + threadIdx.x; idx < n; idx += gridDim.x * blockDim.x) {
    const K k = keys[idx];
    for (int p = 0; p < passes; p++) atomicAdd(&sh[p * 256 + (int)((k >> (first_bit + 8 * p)) & 255)], 1u);
  }
  __syncthreads();
  for (int i = threadIdx.x; i < passes * 256; i += blockDim.x)
    if (sh[i]) atomicAdd(&hist[i], sh[i]);
}

// ---- 2. exclusive scan of each pass's 256 bins (one block per pass) ---------------------------------------------
__device__ __forceinline__ uint32_t block_exclusive_scan_256(uint32_t v, uint32_t* s_warp /* >= 8 */) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  uint32_t incl = v;
#pragma unroll
  for (int off = 1; off < 32; off <<= 1) {
    const uint32_t t = __shfl_up_sync(0xffffffffu, incl, off);
    if (lane >= off) incl += t;
  }
  if (lane == 31) s_warp[warp] = incl;
  __syncthreads();
  uint32_t base = 0;
#pragma unroll
  for (int w = 0; w < kSortWarps; w++)
    if (w < warp) base += s_warp[w];
  __syncthreads();
  return base + incl - v;
}
static __global__ void __launch_bounds__(256) k_sort_scan_hist(uint32_t* __restrict__ hist) {
  __shared__ uint32_t s_warp[kSortWarps];
  uint32_t* h = hist + blockIdx.x * 256;
  const uint32_t v = h[threadIdx.x];
  h[threadIdx.x] = block_exclusive_scan_256(v, s_warp);
}

// ---- 3. one pass -------------------------------------------------------------------------------------------------
// Order inside a tile (defines stability): warp-major, then item, then lane — element (w, i, l) is input index
// tile_base + w*512 + i*32 + l, i.e. plain input order.
template <class K>
struct SortSmem {
  K keys[kSortTile];
  uint32_t vals[kSortTile];
  uint32_t cnt[kSortWarps][256];  // per-warp digit counts, then exclusive offsets of the warp inside its digit
  uint32_t start[256];            // first tile-local position of the digit
  uint32_t base[256];             // global position of the digit's first key of this tile, minus start[]
  uint32_t warp_sums[kSortWarps];
  uint32_t tile;
};

template <class K>
__global__ void __launch_bounds__(kSortBlock, RTF_SORT_MIN_BLOCKS) k_sort_pass(const K* __restrict__ kin, K* __restrict__ kout,
                                                          const uint32_t* __restrict__ vin, uint32_t* __restrict__ vout,
                                                          uint32_t n, int shift, const uint32_t* __restrict__ hist_scan,
                                                          uint32_t* status /* [0] ticket, [256 + tile*256 + d] */) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  SortSmem<K>& sm = *reinterpret_cast<SortSmem<K>*>(smem_raw);
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  if (threadIdx.x == 0) sm.tile = atomicAdd(status, 1u);
#pragma unroll
  for (int w = 0; w < kSortWarps; w++) sm.cnt[w][threadIdx.x] = 0u;
  __syncthreads();
  const uint32_t tile = sm.tile;
  const uint32_t tile_base = tile * (uint32_t)kSortTile;
  const uint32_t count = min((uint32_t)kSortTile, n - tile_base);
  volatile uint32_t* st = status + 256;

  K key[kSortItems];
  uint16_t lrank[kSortItems];
  const uint32_t wbase = tile_base + (uint32_t)warp * 32u * kSortItems + (uint32_t)lane;
  const uint32_t lt_mask = (1u << lane) - 1u;
#pragma unroll
  for (int i = 0; i < kSortItems; i++) {
    const uint32_t idx = wbase + (uint32_t)i * 32u;
    const bool valid = idx < n;
    key[i] = valid ? kin[idx] : (K)0;
    const uint32_t d = valid ? (uint32_t)((key[i] >> shift) & 255) : 0u;
#if RTF_SORT_MATCH
    // invalid lanes (tail of the last tile) get a private digit so they match nobody and count nowhere
    const uint32_t peers = __match_any_sync(0xffffffffu, valid ? d : 256u + (uint32_t)lane);
#else
    // lanes with the same digit, from eight ballots (one per digit bit): match.any serialises over the distinct values
    // of the warp — about 28 of them for random digits
    uint32_t peers = __ballot_sync(0xffffffffu, valid);
#pragma unroll
    for (int b = 0; b < 8; b++) {
      const bool bit = (d >> b) & 1u;
      const uint32_t bal = __ballot_sync(0xffffffffu, bit);
      peers &= bit ? bal : ~bal;
    }
    if (!valid) peers = 1u << lane;
#endif
    const int leader = __ffs((int)peers) - 1;
    uint32_t old = 0;
    if (valid && lane == leader) old = atomicAdd(&sm.cnt[warp][d], (uint32_t)__popc(peers));
    old = __shfl_sync(0xffffffffu, old, leader);
    lrank[i] = (uint16_t)(old + (uint32_t)__popc(peers & lt_mask));
    __syncwarp();
  }
  __syncthreads();
  // thread d owns digit d: warp offsets, tile count, publication, look-back
  {
    const int d = threadIdx.x;
    uint32_t running = 0;
#pragma unroll
    for (int w = 0; w < kSortWarps; w++) {
      const uint32_t c = sm.cnt[w][d];
      sm.cnt[w][d] = running;
      running += c;
    }
    const uint32_t tile_cnt = running;
    st[(size_t)tile * 256 + d] = (tile == 0 ? kSortFlagPrefix : kSortFlagAgg) | tile_cnt;
    const uint32_t start = block_exclusive_scan_256(tile_cnt, sm.warp_sums);
    uint32_t excl = 0;
    if (tile > 0) {
      // look-back over the preceding tiles' counts of digit d, kLook descriptors per round: the loads of a round are
      // independent (one L2 round trip for all of them) — walking back one tile at a time made a pass latency-bound,
      // ~30 dependent reads per tile with some hundred tiles in flight
      constexpr int kLook = 8;
      bool done = false;
      for (int t = (int)tile - 1; !done; t -= kLook) {
        uint32_t v[kLook];
#pragma unroll
        for (int k = 0; k < kLook; k++) v[k] = t - k >= 0 ? st[(size_t)(t - k) * 256 + d] : (uint32_t)(2u << 30);
#pragma unroll
        for (int k = 0; k < kLook; k++) {
          if (done) break;
          while ((v[k] & ~kSortValueMask) == 0u) v[k] = st[(size_t)(t - k) * 256 + d];
          excl += v[k] & kSortValueMask;
          done = (v[k] & kSortFlagPrefix) != 0u;
        }
      }
      st[(size_t)tile * 256 + d] = kSortFlagPrefix | (excl + tile_cnt);
    }
    sm.start[d] = start;
    sm.base[d] = hist_scan[d] + excl - start;
  }
  __syncthreads();
#pragma unroll
  for (int i = 0; i < kSortItems; i++) {
    const uint32_t idx = wbase + (uint32_t)i * 32u;
    if (idx < n) {
      const uint32_t d = (uint32_t)((key[i] >> shift) & 255);
      const uint32_t pos = sm.start[d] + sm.cnt[warp][d] + (uint32_t)lrank[i];
      sm.keys[pos] = key[i];
      sm.vals[pos] = vin ? vin[idx] : idx;  // vin == nullptr: values are the input positions
    }
  }
  __syncthreads();
  for (uint32_t j = threadIdx.x; j < count; j += kSortBlock) {
    const K k = sm.keys[j];
    const uint32_t dst = sm.base[(uint32_t)((k >> shift) & 255)] + j;
    kout[dst] = k;
    vout[dst] = sm.vals[j];
  }
}

// Sorts n pairs by key bits [begin_bit, end_bit).  keys_in / vals_in are preserved (vals_in may be null: the values are
// then 0..n-1).  The result lands in keys_out / vals_out; keys_tmp / vals_tmp are ping-pong scratch of the same size.
// `hist` and `status` are grow-only scratch buffers owned by the caller.
template <class K>
int radix_sort_pairs(rtf_ctx* ctx, const K* keys_in, K* keys_out, K* keys_tmp, const uint32_t* vals_in, uint32_t* vals_out,
                     uint32_t* vals_tmp, uint32_t n, int begin_bit, int end_bit, DevBuf& hist, DevBuf& status) {
  if (n == 0) return RTF_OK;
  if (n > kSortValueMask) {
    set_error("radix sort: more than 2^30 - 1 keys");
    return RTF_ERR_LIMIT;
  }
  cudaStream_t st = ctx->stream;
  const int passes = (end_bit - begin_bit + 7) / 8;
  if (passes < 1 || passes > kSortMaxPasses) return RTF_ERR_INVALID;
  const uint32_t tiles = (n + kSortTile - 1) / kSortTile;
  int rc;
  if ((rc = hist.reserve(4 * 256 * kSortMaxPasses))) return rc;
  if ((rc = status.reserve(4 * (256 + (size_t)tiles * 256)))) return rc;
  RTF_CUDA(cudaMemsetAsync(hist.p, 0, 4 * 256 * (size_t)passes, st));
  const unsigned hist_grid = (unsigned)std::min<size_t>(((size_t)n + 255) / 256, (size_t)ctx->sm_count * 8);
  k_sort_hist<K><<<hist_grid, 256, 0, st>>>(keys_in, n, passes, begin_bit, hist.as<uint32_t>());
  k_sort_scan_hist<<<passes, 256, 0, st>>>(hist.as<uint32_t>());
  ctx->launches += 2;
  const size_t smem = sizeof(SortSmem<K>);
  RTF_CUDA(cudaFuncSetAttribute(k_sort_pass<K>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  const K* kin = keys_in;
  const uint32_t* vin = vals_in;
  for (int p = 0; p < passes; p++) {
    const bool last_parity = ((passes - 1 - p) & 1) == 0;  // the final pass must write keys_out / vals_out
    K* kout = last_parity ? keys_out : keys_tmp;
    uint32_t* vout = last_parity ? vals_out : vals_tmp;
    RTF_CUDA(cudaMemsetAsync(status.p, 0, 4 * (256 + (size_t)tiles * 256), st));
    k_sort_pass<K><<<tiles, kSortBlock, smem, st>>>(kin, kout, vin, vout, n, begin_bit + 8 * p,
                                                    hist.as<uint32_t>() + 256 * p, status.as<uint32_t>());
    ctx->launches++;
    kin = kout, vin = vout;
  }
  RTF_CUDA(cudaGetLastError());
  return RTF_OK;
}

}  // namespace rtf
