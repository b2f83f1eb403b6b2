// a8+a9+a10: CE + Lovasz-Softmax loss, forward and backward (utils/losses.py:19-70,
// LovaszSoftmax/pytorch/lovasz_losses.py:18-30,171-199).
//
//   1. flag rows (mask & label != ignore), count per 256-row tile, scan tile counts      -> N
//   2. compact + softmax + weighted CE + argmax; emit 16 class segments of sort keys |fg - p_c|
//   2b. per class, select what has to be sorted: the foreground rows and the background rows whose error is not below
//      the smallest foreground error.  Everything behind the last foreground element of the sorted order has
//      intersection 0, Jaccard index exactly 1 and therefore gradient exactly 0 (lovasz_losses.py:25-29), so it
//      contributes neither to the loss nor to the gradient and its rank is never needed - bit-identical, and for
//      confident predictions most of the 15 N background keys drop out.  Order-preserving compaction per class.
//   3. per-class segmented LSD radix sort, descending value, stable (ties by ascending index, F14):
//      4 passes of 8 bits over key' = 0x3F800000 - bits(err)   (err in [0,1] => key' in [0, 2^30))
//   4. per class: integer prefix sums of the sorted foreground flags -> Jaccard gradient -> dot
//   5. closed-form backward through |.|, the sort permutation, softmax and CE -> dlogits
// Everything is HBM-bound integer / fp32 work; there is no tensor-core stage.
#include <algorithm>
#include <cstdlib>
#include <mutex>
#include <string.h>

#include "common.cuh"

namespace tdssf {

constexpr int kC = 16;             // classes
constexpr int kRowTile = 256;      // rows per CTA in the flag/compact kernels
constexpr int kSortThreads = 256;
constexpr int kKeysPerThread = 8;
constexpr int kSortTile = kSortThreads * kKeysPerThread;  // 2048 keys per CTA
constexpr int kRadixBits = 8, kBins = 1 << kRadixBits, kPasses = 4;
constexpr int kFusedScanTiles = 256;   // up to this many sort tiles per class the scatter kernel derives its own offsets
static_assert(kBins == kSortThreads, "one thread per radix bin in the histogram / offset code");
constexpr uint32_t kOneBits = 0x3F800000u;

struct CelsWs {
  int* tile_count;      // [n_tiles_rows]
  int* tile_offset;     // [n_tiles_rows]
  int* n_valid;         // [1]
  int* class_count;     // [16]  foreground count per class (G_c)
  uint32_t* fg_keymax;  // [16]  largest sort key (= smallest error) among the foreground rows of the class
  int* seg_count;       // [16]  keys selected for sorting per class
  int* sel_tile;        // [16][row tiles of the compact rows] selected keys per tile, then their exclusive prefix
  double* acc;          // [0]=sum w*nll  [1]=sum w  [2..17]=lovasz loss per class
  int* rows;            // [N]   original row of compact row i
  uint8_t* ylab;        // [N]   label-1
  float* probs;         // [N,16]
  float* dprob;         // [N,16] d lovasz_c / d p_c  (unscaled)
  uint32_t* keys[2];    // [16*N]
  uint32_t* vals[2];    // [16*N]
  uint32_t* hist;       // [16][kBins][sort_tiles]
  int* tile_fg;         // [16][sort_tiles]
  int stride;           // distance between the per-class segments of keys / vals: n_valid (exact mode) or n_rows (device-count mode)
};

static size_t align_up(size_t x) { return (x + 255) & ~(size_t)255; }

static size_t carve(CelsWs* ws, char* base, int64_t n_rows) {
  const int64_t row_tiles = ceil_div(n_rows, kRowTile);
  const int64_t sort_tiles = ceil_div(n_rows, kSortTile);
  size_t off = 0;
  auto take = [&](size_t bytes) { char* p = base ? base + off : nullptr; off += align_up(bytes); return p; };
  char* p;
  p = take(row_tiles * sizeof(int)); if (ws) ws->tile_count = (int*)p;
  p = take(row_tiles * sizeof(int)); if (ws) ws->tile_offset = (int*)p;
  p = take(sizeof(int)); if (ws) ws->n_valid = (int*)p;              // } zeroed by one memset at the start of a call:
  p = take(kC * sizeof(int)); if (ws) ws->class_count = (int*)p;     // } keep these five together and in this order
  p = take(kC * sizeof(uint32_t)); if (ws) ws->fg_keymax = (uint32_t*)p;
  p = take(kC * sizeof(int)); if (ws) ws->seg_count = (int*)p;
  p = take(18 * sizeof(double)); if (ws) ws->acc = (double*)p;
  p = take((size_t)kC * row_tiles * sizeof(int)); if (ws) ws->sel_tile = (int*)p;
  p = take(n_rows * sizeof(int)); if (ws) ws->rows = (int*)p;
  p = take(n_rows); if (ws) ws->ylab = (uint8_t*)p;
  p = take(n_rows * kC * sizeof(float)); if (ws) ws->probs = (float*)p;
  p = take(n_rows * kC * sizeof(float)); if (ws) ws->dprob = (float*)p;
  for (int i = 0; i < 2; ++i) {
    p = take(n_rows * kC * sizeof(uint32_t)); if (ws) ws->keys[i] = (uint32_t*)p;
    p = take(n_rows * kC * sizeof(uint32_t)); if (ws) ws->vals[i] = (uint32_t*)p;
  }
  p = take((size_t)kC * kBins * sort_tiles * sizeof(uint32_t)); if (ws) ws->hist = (uint32_t*)p;
  p = take((size_t)kC * sort_tiles * sizeof(int)); if (ws) ws->tile_fg = (int*)p;
  return off + 256;
}

// ---- block-wide helpers (256 threads) -----------------------------------------------------------
__device__ __forceinline__ int warp_incl_scan(int v) {
  const int lane = threadIdx.x & 31;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    int t = __shfl_up_sync(0xffffffffu, v, o);
    if (lane >= o) v += t;
  }
  return v;
}
// exclusive scan of one int per thread over a 256-thread block; *total = block sum
__device__ __forceinline__ int block_excl_scan256(int v, int* s_warp /*[9]*/, int* total) {
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  const int incl = warp_incl_scan(v);
  if (lane == 31) s_warp[w] = incl;
  __syncthreads();
  if (w == 0) {
    int x = lane < 8 ? s_warp[lane] : 0;
    int xi = warp_incl_scan(x);
    if (lane < 8) s_warp[lane] = xi - x;
    if (lane == 7) s_warp[8] = xi;
  }
  __syncthreads();
  const int r = s_warp[w] + incl - v;
  *total = s_warp[8];
  __syncthreads();
  return r;
}

__device__ __forceinline__ bool row_is_valid(const int64_t* labels, const uint8_t* mask, int64_t r, int ignore) {
  if (mask && !mask[r]) return false;
  return labels[r] != (int64_t)ignore;
}

// sort key of |fg - p|: ascending key = descending error; err in [0,1] => key in [0, 0x3F800000]
__device__ __forceinline__ uint32_t sort_key(bool fg, float p) {
  const float err = fabsf((fg ? 1.f : 0.f) - p);
  return kOneBits - min(__float_as_uint(err), kOneBits);
}

// ---- 1. flags ----------------------------------------------------------------------------------
__global__ void __launch_bounds__(kRowTile)
cels_count_kernel(const int64_t* __restrict__ labels, const uint8_t* __restrict__ mask, int64_t n_rows,
                  int ignore, int* __restrict__ tile_count) {
  const int64_t r = (int64_t)blockIdx.x * kRowTile + threadIdx.x;
  const int f = (r < n_rows) && row_is_valid(labels, mask, r, ignore);
  const int c = __syncthreads_count(f);
  if (threadIdx.x == 0) tile_count[blockIdx.x] = c;
}

// single CTA: exclusive scan of `n` ints (n up to millions: chunked), total -> *out_total
__global__ void __launch_bounds__(256)
scan_ints_kernel(const int* __restrict__ in, int* __restrict__ out, int64_t n, int* __restrict__ out_total) {
  __shared__ int s_warp[9];
  int carry = 0;
  for (int64_t base = 0; base < n; base += 256) {
    const int64_t i = base + threadIdx.x;
    const int v = i < n ? in[i] : 0;
    int tot;
    const int e = block_excl_scan256(v, s_warp, &tot);
    if (i < n) out[i] = carry + e;
    carry += tot;
  }
  if (threadIdx.x == 0 && out_total) *out_total = carry;
}

// ---- 2. compact + softmax + CE + keys ------------------------------------------------------------
__global__ void __launch_bounds__(kRowTile)
cels_compact_kernel(const float* __restrict__ logits, const int64_t* __restrict__ labels,
                    const uint8_t* __restrict__ mask, int64_t n_rows, int ignore,
                    const float* __restrict__ class_w, CelsWs ws, int64_t* __restrict__ preds,
                    int64_t* __restrict__ gt) {
  __shared__ int s_warp[9];
  __shared__ int s_cls[kC];
  __shared__ uint32_t s_kmax[kC];
  __shared__ double s_red[2][8];
  if (threadIdx.x < kC) { s_cls[threadIdx.x] = 0; s_kmax[threadIdx.x] = 0u; }
  const int64_t r = (int64_t)blockIdx.x * kRowTile + threadIdx.x;
  const int f = (r < n_rows) && row_is_valid(labels, mask, r, ignore);
  int tot;
  const int local = block_excl_scan256(f, s_warp, &tot);  // also orders the s_cls init
  const int N = *ws.n_valid;
  double wnll = 0.0, wsum = 0.0;
  if (f) {
    const int i = ws.tile_offset[blockIdx.x] + local;
    float x[kC];
    const float4* src = reinterpret_cast<const float4*>(logits + r * kC);
#pragma unroll
    for (int q = 0; q < 4; ++q) {
      float4 v = __ldg(src + q);
      x[4 * q] = v.x; x[4 * q + 1] = v.y; x[4 * q + 2] = v.z; x[4 * q + 3] = v.w;
    }
    const int y = (int)labels[r] - 1;  // 0..15 (ignore_index rows were dropped)
    float m = x[0];
#pragma unroll
    for (int c = 1; c < kC; ++c) m = fmaxf(m, x[c]);
    float e[kC], s = 0.f;
#pragma unroll
    for (int c = 0; c < kC; ++c) { e[c] = expf(x[c] - m); s += e[c]; }
    float p[kC];
    int am = 0;
    float pm = -1.f;
#pragma unroll
    for (int c = 0; c < kC; ++c) {
      p[c] = __fdiv_rn(e[c], s);
      if (p[c] > pm) { pm = p[c]; am = c; }  // first maximum wins, like torch.argmax
    }
    float xy = x[0];
#pragma unroll
    for (int c = 1; c < kC; ++c) xy = (c == y) ? x[c] : xy;
    const bool y_ok = (y >= 0 && y < kC);
    const float w = y_ok ? __ldg(class_w + y) : 0.f;
    const float nll = logf(s) - (xy - m);
    wnll = (double)w * (double)nll;
    wsum = (double)w;
    preds[i] = am + 1;
    gt[i] = y + 1;
    ws.rows[i] = (int)r;
    ws.ylab[i] = (uint8_t)y;
    float4* pd = reinterpret_cast<float4*>(ws.probs + (int64_t)i * kC);
#pragma unroll
    for (int q = 0; q < 4; ++q) pd[q] = make_float4(p[4 * q], p[4 * q + 1], p[4 * q + 2], p[4 * q + 3]);
    if (y_ok) {
      float py = p[0];
#pragma unroll
      for (int c = 1; c < kC; ++c) py = (c == y) ? p[c] : py;
      atomicAdd(&s_cls[y], 1);
      atomicMax(&s_kmax[y], sort_key(true, py));
    }
  }
  // block reduction of the CE partial sums
  const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    wnll += __shfl_xor_sync(0xffffffffu, wnll, o);
    wsum += __shfl_xor_sync(0xffffffffu, wsum, o);
  }
  if (lane == 0) { s_red[0][wid] = wnll; s_red[1][wid] = wsum; }
  __syncthreads();
  if (threadIdx.x == 0) {
    double a = 0, b = 0;
    for (int i = 0; i < 8; ++i) { a += s_red[0][i]; b += s_red[1][i]; }
    if (b != 0.0 || a != 0.0) { atomicAdd(ws.acc + 0, a); atomicAdd(ws.acc + 1, b); }
  }
  if (threadIdx.x < kC && s_cls[threadIdx.x]) {
    atomicAdd(ws.class_count + threadIdx.x, s_cls[threadIdx.x]);
    atomicMax(ws.fg_keymax + threadIdx.x, s_kmax[threadIdx.x]);
  }
}

// ---- 2b. selection of the keys that need a rank ---------------------------------------------------------------------
// One thread per compact row, 16 ballots per warp.  kEmit == false: selected keys per (class, tile) -> sel_tile;
// kEmit == true: sel_tile holds the exclusive prefix over tiles, keys/values are written in row order (the stable sort's
// tie-break is the initial order).
template <bool kEmit>
__global__ void __launch_bounds__(kRowTile)
lovasz_select_kernel(CelsWs ws, int tiles) {
  __shared__ int s_wc[kRowTile / 32][kC];
  const int N = *ws.n_valid;
  const int i = blockIdx.x * kRowTile + threadIdx.x;
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  float p[kC];
  int y = -1;
  if (i < N) {
    const float4* ps = reinterpret_cast<const float4*>(ws.probs + (int64_t)i * kC);
#pragma unroll
    for (int q = 0; q < 4; ++q) {
      const float4 v = ps[q];
      p[4 * q] = v.x; p[4 * q + 1] = v.y; p[4 * q + 2] = v.z; p[4 * q + 3] = v.w;
    }
    y = ws.ylab[i];
  }
  uint32_t key[kC], sel = 0;   // bit c of sel: this row's key of class c is sorted
  int rank[kC];
#pragma unroll
  for (int c = 0; c < kC; ++c) {
    key[c] = i < N ? sort_key(c == y, p[c]) : 0u;
    const bool inc = i < N && ws.class_count[c] > 0 && (c == y || key[c] <= ws.fg_keymax[c]);
    const uint32_t b = __ballot_sync(0xffffffffu, inc);
    sel |= (inc ? 1u : 0u) << c;
    rank[c] = __popc(b & ((1u << lane) - 1));
    if (lane == 0) s_wc[w][c] = __popc(b);
  }
  __syncthreads();
  if (!kEmit) {
    if (threadIdx.x < kC) {
      int tot = 0;
#pragma unroll
      for (int ww = 0; ww < kRowTile / 32; ++ww) tot += s_wc[ww][threadIdx.x];
      ws.sel_tile[threadIdx.x * tiles + blockIdx.x] = tot;
    }
    return;
  }
#pragma unroll
  for (int c = 0; c < kC; ++c) {
    if ((sel >> c) & 1u) {
      int before = ws.sel_tile[c * tiles + blockIdx.x];
      for (int ww = 0; ww < w; ++ww) before += s_wc[ww][c];
      const int64_t pos = (int64_t)c * ws.stride + before + rank[c];
      ws.keys[0][pos] = key[c];
      ws.vals[0][pos] = (uint32_t)i;
    }
  }
}

// one CTA per class: exclusive scan of the per-tile selection counts, total -> seg_count
__global__ void __launch_bounds__(256)
lovasz_segscan_kernel(CelsWs ws, int tiles) {
  const int seg = blockIdx.x;
  __shared__ int s_warp[9];
  int* t = ws.sel_tile + seg * tiles;
  int carry = 0;
  for (int base = 0; base < tiles; base += 256) {
    const int i = base + threadIdx.x;
    const int v = i < tiles ? t[i] : 0;
    int tot;
    const int e = block_excl_scan256(v, s_warp, &tot);
    if (i < tiles) t[i] = carry + e;
    carry += tot;
  }
  if (threadIdx.x == 0) ws.seg_count[seg] = carry;
}

// ---- 3. segmented LSD radix sort ------------------------------------------------------------------
// grid = (sort_tiles, 16 segments).  hist layout [seg][bin][tile] so that one exclusive scan per
// segment yields the global offset of every (bin, tile) pair.
__global__ void __launch_bounds__(kSortThreads)
sort_hist_kernel(const uint32_t* __restrict__ keys, int Nstride, const int* __restrict__ seg_count, int shift, int tiles,
                 const int* __restrict__ class_count, uint32_t* __restrict__ hist, bool tile_major) {
  const int seg = blockIdx.y, tile = blockIdx.x;
  const int N = seg_count[seg];   // keys of this class that are sorted; the grid is sized for all valid rows
  if (class_count[seg] == 0 || tile * kSortTile >= N) return;  // absent class: skipped by classes='present'
  __shared__ uint32_t s_h[kBins];
  s_h[threadIdx.x] = 0;
  __syncthreads();
  const uint32_t* k = keys + (int64_t)seg * Nstride;
  const int base = tile * kSortTile;
#pragma unroll
  for (int j = 0; j < kKeysPerThread; ++j) {
    const int idx = base + j * kSortThreads + threadIdx.x;
    if (idx < N) atomicAdd(&s_h[(k[idx] >> shift) & (kBins - 1)], 1u);
  }
  __syncthreads();
  if (tile_major) hist[((int64_t)seg * tiles + tile) * kBins + threadIdx.x] = s_h[threadIdx.x];   // raw counts, see sort_scatter_kernel
  else hist[((int64_t)seg * kBins + threadIdx.x) * tiles + tile] = s_h[threadIdx.x];
}

// one CTA per segment: in-place exclusive scan over kBins*tiles counters, 8 per thread per step
__global__ void __launch_bounds__(256)
sort_scan_kernel(uint32_t* __restrict__ hist, int tiles, const int* __restrict__ class_count,
                 const int* __restrict__ seg_count) {
  const int seg = blockIdx.x;
  if (class_count[seg] == 0) return;
  __shared__ int s_warp[9];
  uint32_t* h = hist + (int64_t)seg * kBins * tiles;
  const int64_t n = (int64_t)kBins * tiles;
  const int tiles_act = (seg_count[seg] + kSortTile - 1) / kSortTile;  // [bin][tile] layout: later tiles were not written
  int carry = 0;
  for (int64_t base = 0; base < n; base += 2048) {
    const int64_t i0 = base + threadIdx.x * 8;
    uint32_t v[8];
    int sum = 0;
#pragma unroll
    for (int j = 0; j < 8; ++j) { v[j] = (i0 + j < n && (int)((i0 + j) % tiles) < tiles_act) ? h[i0 + j] : 0u; sum += (int)v[j]; }
    int tot;
    int run = carry + block_excl_scan256(sum, s_warp, &tot);
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      if (i0 + j < n) h[i0 + j] = (uint32_t)run;
      run += (int)v[j];
    }
    carry += tot;
  }
}

#ifndef TDSSF_SCATTER_MINB
#define TDSSF_SCATTER_MINB 5  // 48 registers: five CTAs per SM (58 registers and four CTAs without the bound: 6 % slower)
#endif
__global__ void __launch_bounds__(kSortThreads, TDSSF_SCATTER_MINB)
sort_scatter_kernel(const uint32_t* __restrict__ keys_in, const uint32_t* __restrict__ vals_in,
                    uint32_t* __restrict__ keys_out, uint32_t* __restrict__ vals_out, int Nstride,
                    const int* __restrict__ seg_count, int shift, int tiles, const int* __restrict__ class_count,
                    const uint32_t* __restrict__ hist, bool tile_major) {
  const int seg = blockIdx.y, tile = blockIdx.x;
  const int N = seg_count[seg];
  if (class_count[seg] == 0 || tile * kSortTile >= N) return;
  const int tiles_act = (N + kSortTile - 1) / kSortTile;  // tiles that hold keys (the only histograms written)
  constexpr int kWarps = kSortThreads / 32;
  __shared__ uint32_t s_cnt[kWarps][kBins];   // per-warp digit counters, then exclusive prefix over warps
  __shared__ uint32_t s_base[kBins];          // global offset of (bin, tile)
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  for (int i = threadIdx.x; i < kWarps * kBins; i += kSortThreads) (&s_cnt[0][0])[i] = 0;
  // warp w owns 256 consecutive keys of the tile, processed as 8 rounds of 32 consecutive keys so
  // that (round, lane) order == key order  => the ranking below is stable.  All 16 loads of the thread are issued
  // before anything waits (one round trip to L2 instead of eight), and they overlap the offset computation below.
  const int64_t so = (int64_t)seg * Nstride;
  const int wbase = tile * kSortTile + w * (32 * kKeysPerThread);
  uint32_t key[kKeysPerThread], val[kKeysPerThread], rank[kKeysPerThread];
#pragma unroll
  for (int j = 0; j < kKeysPerThread; ++j) {
    const int idx = wbase + j * 32 + lane;
    const bool in = idx < N;
    key[j] = in ? __ldg(keys_in + so + idx) : 0xffffffffu;
    val[j] = in ? __ldg(vals_in + so + idx) : 0u;
  }
  if (tile_major) {
    // up to kFusedScanTiles tiles: no scan launch.  hist holds the raw [seg][tile][bin] counts and every CTA derives the
    // offset of its (bin, tile) pair itself: keys of smaller bins in all tiles + keys of its bin in earlier tiles
    // (`tiles` coalesced, independent loads per thread from L2, 16 in flight at a time, one block scan).
    __shared__ int s_warp[9];
    const uint32_t* h = hist + (int64_t)seg * tiles * kBins + threadIdx.x;
    uint32_t tot = 0, before = 0;
    int t = 0;
    for (; t + 16 <= tiles_act; t += 16) {
      uint32_t v[16];
#pragma unroll
      for (int u = 0; u < 16; ++u) v[u] = __ldg(h + (int64_t)(t + u) * kBins);
#pragma unroll
      for (int u = 0; u < 16; ++u) { tot += v[u]; before += (t + u) < tile ? v[u] : 0u; }
    }
    for (; t < tiles_act; ++t) {
      const uint32_t v = __ldg(h + (int64_t)t * kBins);
      tot += v;
      before += t < tile ? v : 0u;
    }
    int all;
    s_base[threadIdx.x] = (uint32_t)block_excl_scan256((int)tot, s_warp, &all) + before;
  } else {
    s_base[threadIdx.x] = hist[((int64_t)seg * kBins + threadIdx.x) * tiles + tile];
  }
  __syncthreads();

  // The eight match operations of a thread do not depend on the counters, so they are all issued first (their latency
  // overlaps); out-of-range lanes get a digit no key can have.  rank[j] holds the peer mask until the second loop.
#pragma unroll
  for (int j = 0; j < kKeysPerThread; ++j) {
    const bool in = wbase + j * 32 + lane < N;
    const uint32_t d = (key[j] >> shift) & (kBins - 1);
    rank[j] = __match_any_sync(0xffffffffu, in ? d : (0x100u | (uint32_t)lane));
  }
#pragma unroll
  for (int j = 0; j < kKeysPerThread; ++j) {
    const bool in = wbase + j * 32 + lane < N;
    const uint32_t d = (key[j] >> shift) & (kBins - 1);
    const uint32_t peers = rank[j];
    uint32_t r = 0;
    if (in) {
      const int leader = __ffs(peers) - 1;
      uint32_t old = 0;
      if (lane == leader) { old = s_cnt[w][d]; s_cnt[w][d] = old + __popc(peers); }
      old = __shfl_sync(peers, old, leader);
      r = old + __popc(peers & ((1u << lane) - 1));
    }
    rank[j] = r;
    __syncwarp();
  }
  __syncthreads();
  {  // exclusive prefix over warps for digit = threadIdx.x
    uint32_t run = 0;
#pragma unroll
    for (int ww = 0; ww < kWarps; ++ww) {
      const uint32_t c = s_cnt[ww][threadIdx.x];
      s_cnt[ww][threadIdx.x] = run;
      run += c;
    }
  }
  __syncthreads();
  // (measured: ordering the tile by digit in shared memory first, so that a warp stores runs of equal digits, made the
  // pass slower - 0.399 against 0.365 ms for the whole loss at 4 x 34 720 points)
#pragma unroll
  for (int j = 0; j < kKeysPerThread; ++j) {
    const int idx = wbase + j * 32 + lane;
    if (idx < N) {
      const uint32_t d = (key[j] >> shift) & (kBins - 1);
      const uint32_t pos = s_base[d] + s_cnt[w][d] + rank[j];
      keys_out[so + pos] = key[j];
      vals_out[so + pos] = val[j];
    }
  }
}

// ---- 4. Jaccard gradient + dot --------------------------------------------------------------------
__global__ void __launch_bounds__(kSortThreads)
lovasz_fgcount_kernel(const uint32_t* __restrict__ vals, const uint8_t* __restrict__ ylab, int Nstride,
                      const int* __restrict__ seg_count, int tiles, const int* __restrict__ class_count,
                      int* __restrict__ tile_fg) {
  const int seg = blockIdx.y, tile = blockIdx.x;
  if (class_count[seg] == 0) return;
  const int N = seg_count[seg];   // (tiles behind the selected keys write a zero count: the tile scan reads all of them)
  int c = 0;
  const int base = tile * kSortTile + threadIdx.x * kKeysPerThread;
#pragma unroll
  for (int j = 0; j < kKeysPerThread; ++j) {
    const int idx = base + j;
    if (idx < N) c += (ylab[vals[(int64_t)seg * Nstride + idx]] == seg);
  }
  __shared__ int s_warp[9];
  int tot;
  block_excl_scan256(c, s_warp, &tot);
  if (threadIdx.x == 0) tile_fg[seg * tiles + tile] = tot;
}

__global__ void __launch_bounds__(256)
lovasz_tilescan_kernel(int* __restrict__ tile_fg, int tiles, const int* __restrict__ class_count) {
  const int seg = blockIdx.x;
  if (class_count[seg] == 0) return;
  __shared__ int s_warp[9];
  int* t = tile_fg + seg * tiles;
  int carry = 0;
  for (int base = 0; base < tiles; base += 256) {
    const int i = base + threadIdx.x;
    const int v = i < tiles ? t[i] : 0;
    int tot;
    const int e = block_excl_scan256(v, s_warp, &tot);
    if (i < tiles) t[i] = carry + e;
    carry += tot;
  }
}

__device__ __forceinline__ float jaccard_at(int G, int k /*0-based position*/, int cf /*inclusive fg count*/) {
  // lovasz_grad: 1 - (gts - cumsum(fg)) / (gts + cumsum(1 - fg)); every operand is an exact fp32 integer
  const float inter = (float)(G - cf);
  const float uni = (float)(G + (k + 1 - cf));
  return __fsub_rn(1.f, __fdiv_rn(inter, uni));
}

__global__ void __launch_bounds__(kSortThreads)
lovasz_grad_dot_kernel(const uint32_t* __restrict__ keys, const uint32_t* __restrict__ vals,
                       const uint8_t* __restrict__ ylab, int Nstride, const int* __restrict__ seg_count, int tiles,
                       const int* __restrict__ class_count, const int* __restrict__ tile_fg,
                       float* __restrict__ dprob, double* __restrict__ acc) {
  const int seg = blockIdx.y, tile = blockIdx.x;
  const int G = class_count[seg];
  const int N = seg_count[seg];
  if (G == 0 || tile * kSortTile >= N) return;
  __shared__ int s_warp[9];
  __shared__ double s_red[8];
  const int base = tile * kSortTile + threadIdx.x * kKeysPerThread;
  uint32_t v[kKeysPerThread];
  int fg[kKeysPerThread], c = 0;
#pragma unroll
  for (int j = 0; j < kKeysPerThread; ++j) {
    const int idx = base + j;
    v[j] = idx < N ? vals[(int64_t)seg * Nstride + idx] : 0u;
    fg[j] = idx < N ? (ylab[v[j]] == seg) : 0;
    c += fg[j];
  }
  int tot;
  int cf = tile_fg[seg * tiles + tile] + block_excl_scan256(c, s_warp, &tot);  // fg count before my first key
  double dot = 0.0;
#pragma unroll
  for (int j = 0; j < kKeysPerThread; ++j) {
    const int idx = base + j;
    if (idx < N) {
      const int cf_prev = cf;
      cf += fg[j];
      const float jac = jaccard_at(G, idx, cf);
      const float g = idx == 0 ? jac : __fsub_rn(jac, jaccard_at(G, idx - 1, cf_prev));
      const float err = __uint_as_float(kOneBits - keys[(int64_t)seg * Nstride + idx]);
      dot += (double)err * (double)g;
      // d|fg - p| / dp = -sign(fg - p), sign(0) = 0
      const float sgn = err == 0.f ? 0.f : (fg[j] ? -1.f : 1.f);
      dprob[(int64_t)v[j] * kC + seg] = g * sgn;
    }
  }
  const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) dot += __shfl_xor_sync(0xffffffffu, dot, o);
  if (lane == 0) s_red[wid] = dot;
  __syncthreads();
  if (threadIdx.x == 0) {
    double a = 0;
    for (int i = 0; i < 8; ++i) a += s_red[i];
    atomicAdd(acc + 2 + seg, a);
  }
}

// ---- 5. scalars and backward ----------------------------------------------------------------------
__global__ void cels_finalize_kernel(CelsWs ws, float lambda, float* __restrict__ out) {
  int n_present = 0;
  double lov = 0.0;
  for (int c = 0; c < kC; ++c)
    if (ws.class_count[c] > 0) { n_present++; lov += ws.acc[2 + c]; }
  // mean() helper of lovasz_losses.py:232-250: acc / n  (n == 1 returns acc itself)
  const float lovasz = n_present > 0 ? (float)(lov / n_present) : 0.f;
  const float ce = (float)(ws.acc[0] / ws.acc[1]);  // 0/0 = NaN on an empty selection, like torch
  out[0] = ce + lambda * lovasz;
  out[1] = ce;
  out[2] = lovasz;
  out[3] = (float)n_present;
}

__global__ void __launch_bounds__(256)
cels_backward_kernel(CelsWs ws, const float* __restrict__ class_w, float a_ce, float a_lov, const float* __restrict__ coef,
                     float* __restrict__ dlogits) {
  // dlogits = a_ce * d ce / d logits + a_lov * d lovasz / d logits; the forward call uses (1, lambda) = d total; `coef`
  // (device, optional) overrides both with the upstream gradients of a backward pass through ce / lovasz themselves
  const int N = *ws.n_valid;
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= N) return;
  if (coef) { a_ce = __ldg(coef); a_lov = __ldg(coef + 1); }
  int n_present = 0;
#pragma unroll
  for (int c = 0; c < kC; ++c) n_present += ws.class_count[c] > 0;
  const float lscale = n_present > 0 ? a_lov / (float)n_present : 0.f;
  const float wsum = (float)ws.acc[1];
  const int y = ws.ylab[i];
  const float wy = y < kC ? a_ce * (__ldg(class_w + y) / wsum) : 0.f;
  float p[kC], dp[kC], s = 0.f;
  const float4* ps = reinterpret_cast<const float4*>(ws.probs + (int64_t)i * kC);
  const float4* ds = reinterpret_cast<const float4*>(ws.dprob + (int64_t)i * kC);
#pragma unroll
  for (int q = 0; q < 4; ++q) {
    float4 a = ps[q], b = ds[q];
    p[4 * q] = a.x; p[4 * q + 1] = a.y; p[4 * q + 2] = a.z; p[4 * q + 3] = a.w;
    dp[4 * q] = b.x * lscale; dp[4 * q + 1] = b.y * lscale; dp[4 * q + 2] = b.z * lscale; dp[4 * q + 3] = b.w * lscale;
  }
#pragma unroll
  for (int c = 0; c < kC; ++c) s += dp[c] * p[c];
  float g[kC];
#pragma unroll
  for (int c = 0; c < kC; ++c) g[c] = p[c] * (dp[c] - s) + wy * (p[c] - (c == y ? 1.f : 0.f));
  float4* out = reinterpret_cast<float4*>(dlogits + (int64_t)ws.rows[i] * kC);
#pragma unroll
  for (int q = 0; q < 4; ++q) out[q] = make_float4(g[4 * q], g[4 * q + 1], g[4 * q + 2], g[4 * q + 3]);
}

}  // namespace tdssf

extern "C" size_t tdssf_cels_workspace_bytes(int64_t n_rows) {
  if (n_rows < 0) n_rows = 0;
  return tdssf::carve(nullptr, nullptr, std::max<int64_t>(n_rows, 1));
}

namespace tdssf {

struct CelsArgs {
  const float* logits; const int64_t* labels; const uint8_t* mask; int64_t n_rows; const float* class_w; float lambda; int ignore;
  float* out_scalars; int64_t* preds; int64_t* gt; float* dlogits; void* workspace;
  bool operator==(const CelsArgs& o) const { return memcmp(this, &o, sizeof(CelsArgs)) == 0; }
};

// stage 1: valid-row flags per tile and their exclusive scan; *ws.n_valid = number of rows that survive mask & label != ignore
static int cels_stage_count(const CelsArgs& a, const CelsWs& ws, cudaStream_t s) {
  // n_valid, class_count, fg_keymax, seg_count and acc are carved next to each other: one memset
  TDSSF_CUDA(cudaMemsetAsync(ws.n_valid, 0, (size_t)(reinterpret_cast<char*>(ws.acc + 18) - reinterpret_cast<char*>(ws.n_valid)), s));
  if (a.n_rows > 0) {
    const unsigned row_tiles = (unsigned)ceil_div(a.n_rows, kRowTile);
    cels_count_kernel<<<row_tiles, kRowTile, 0, s>>>(a.labels, a.mask, a.n_rows, a.ignore, ws.tile_count);
    TDSSF_LAUNCH_CHECK();
    scan_ints_kernel<<<1, 256, 0, s>>>(ws.tile_count, ws.tile_offset, row_tiles, ws.n_valid);
    TDSSF_LAUNCH_CHECK();
  }
  return TDSSF_OK;
}

// stage 2: everything after the count.  `bound` = number of compact rows the grids and the segment stride are sized for:
// the exact n_valid when the host has read it (tdssf_cels_fwd_bwd), or n_rows when it has not (device-count mode: every
// kernel takes its real bounds from the counts in the workspace, the surplus CTAs return at once) - the same results.
static int cels_stage_main(const CelsArgs& a, CelsWs ws, int64_t bound, cudaStream_t s) {
  if (a.dlogits && a.n_rows) TDSSF_CUDA(cudaMemsetAsync(a.dlogits, 0, (size_t)a.n_rows * kC * sizeof(float), s));
  if (bound > 0) {
    const int N = (int)bound;
    ws.stride = N;
    const unsigned row_tiles = (unsigned)ceil_div(a.n_rows, kRowTile);
    cels_compact_kernel<<<row_tiles, kRowTile, 0, s>>>(a.logits, a.labels, a.mask, a.n_rows, a.ignore, a.class_w, ws, a.preds, a.gt);
    TDSSF_LAUNCH_CHECK();
    // TDSSF_CELS_SELECT=all (tests): every key of a present class is sorted, as the reference does
    const char* sel_mode = getenv("TDSSF_CELS_SELECT");
    if (sel_mode && sel_mode[0] == 'a') TDSSF_CUDA(cudaMemsetAsync(ws.fg_keymax, 0xff, kC * sizeof(uint32_t), s));
    const int sel_tiles = (int)ceil_div(N, kRowTile);
    lovasz_select_kernel<false><<<sel_tiles, kRowTile, 0, s>>>(ws, sel_tiles);
    TDSSF_LAUNCH_CHECK();
    lovasz_segscan_kernel<<<kC, 256, 0, s>>>(ws, sel_tiles);
    TDSSF_LAUNCH_CHECK();
    lovasz_select_kernel<true><<<sel_tiles, kRowTile, 0, s>>>(ws, sel_tiles);
    TDSSF_LAUNCH_CHECK();
    const int tiles = (int)ceil_div(N, kSortTile);   // upper bound: the kernels read the selected count per class
    const dim3 sgrid((unsigned)tiles, kC);
    int cur = 0;
    // beyond kFusedScanTiles the per-CTA prefix reads would grow quadratically (TDSSF_CELS_SCAN=separate: tests)
    const char* scan_mode = getenv("TDSSF_CELS_SCAN");
    const bool fused_scan = tiles <= kFusedScanTiles && !(scan_mode && scan_mode[0] == 's');
    for (int pass = 0; pass < kPasses; ++pass) {
      const int shift = pass * kRadixBits;
      sort_hist_kernel<<<sgrid, kSortThreads, 0, s>>>(ws.keys[cur], N, ws.seg_count, shift, tiles, ws.class_count, ws.hist, fused_scan);
      TDSSF_LAUNCH_CHECK();
      if (!fused_scan) {
        sort_scan_kernel<<<kC, 256, 0, s>>>(ws.hist, tiles, ws.class_count, ws.seg_count);
        TDSSF_LAUNCH_CHECK();
      }
      sort_scatter_kernel<<<sgrid, kSortThreads, 0, s>>>(ws.keys[cur], ws.vals[cur], ws.keys[cur ^ 1], ws.vals[cur ^ 1],
                                                         N, ws.seg_count, shift, tiles, ws.class_count, ws.hist, fused_scan);
      TDSSF_LAUNCH_CHECK();
      cur ^= 1;
    }
    TDSSF_CUDA(cudaMemsetAsync(ws.dprob, 0, (size_t)N * kC * sizeof(float), s));
    lovasz_fgcount_kernel<<<sgrid, kSortThreads, 0, s>>>(ws.vals[cur], ws.ylab, N, ws.seg_count, tiles, ws.class_count, ws.tile_fg);
    TDSSF_LAUNCH_CHECK();
    lovasz_tilescan_kernel<<<kC, 256, 0, s>>>(ws.tile_fg, tiles, ws.class_count);
    TDSSF_LAUNCH_CHECK();
    lovasz_grad_dot_kernel<<<sgrid, kSortThreads, 0, s>>>(ws.keys[cur], ws.vals[cur], ws.ylab, N, ws.seg_count, tiles, ws.class_count,
                                                          ws.tile_fg, ws.dprob, ws.acc);
    TDSSF_LAUNCH_CHECK();
    if (a.dlogits) {
      cels_backward_kernel<<<(unsigned)ceil_div(N, 256), 256, 0, s>>>(ws, a.class_w, 1.f, a.lambda, nullptr, a.dlogits);
      TDSSF_LAUNCH_CHECK();
    }
  }
  cels_finalize_kernel<<<1, 1, 0, s>>>(ws, a.lambda, a.out_scalars);
  TDSSF_LAUNCH_CHECK();
  return TDSSF_OK;
}

static int cels_check(const CelsArgs& a, size_t workspace_bytes, const char* who) {
  TDSSF_REQUIRE(a.n_rows >= 0 && a.n_rows < (int64_t)1 << 27, "%s: n_rows=%lld out of range", who, (long long)a.n_rows);
  TDSSF_REQUIRE(a.out_scalars && a.class_w && a.workspace, "%s: NULL pointer", who);
  TDSSF_REQUIRE(workspace_bytes >= tdssf_cels_workspace_bytes(a.n_rows), "%s: workspace too small", who);
  if (a.n_rows > 0) {
    TDSSF_REQUIRE(a.logits && a.labels && a.preds && a.gt, "%s: NULL input/output", who);
    TDSSF_REQUIRE(((uintptr_t)a.logits & 15) == 0 && (!a.dlogits || ((uintptr_t)a.dlogits & 15) == 0), "%s: logits/dlogits must be 16-byte aligned", who);
  }
  return TDSSF_OK;
}

// ---- device-count mode as a CUDA graph --------------------------------------------------------------------------
// The launch sequence of stage 2 depends only on the argument values (pointers, n_rows), not on the data: when a call
// repeats the arguments of an earlier one (a training loop: the caching allocator hands out the same blocks every
// step, or alternates between a few) the sequence is captured once into a graph and replayed - one launch instead of
// about twenty-five.
struct CelsGraph { CelsArgs key; int device; cudaGraphExec_t exec; unsigned long long used; };
constexpr int kGraphSlots = 8;
static CelsGraph g_graphs[kGraphSlots];
static CelsArgs g_seen[kGraphSlots];   // argument sets of recent plain calls (a loop whose buffers alternate between two or
static int g_seen_dev[kGraphSlots];    // three allocator blocks repeats with that period, not every call)
static int g_seen_next = 0;
static unsigned long long g_tick = 0;
static long long g_stats[3] = {0, 0, 0};   // replays, captures, plain launch sequences
static std::mutex g_graph_mu;

static int cels_stage_main_graphed(const CelsArgs& a, const CelsWs& ws, cudaStream_t s) {
  const char* env = getenv("TDSSF_CELS_GRAPH");
  if (env && env[0] == '0') return cels_stage_main(a, ws, a.n_rows, s);
  std::lock_guard<std::mutex> lock(g_graph_mu);
  const int dev = current_device();
  ++g_tick;
  for (auto& g : g_graphs)
    if (g.exec && g.device == dev && g.key == a) {
      g.used = g_tick;
      ++g_stats[0];
      TDSSF_CUDA(cudaGraphLaunch(g.exec, s));
      return TDSSF_OK;
    }
  bool repeat = false;
  for (int i = 0; i < kGraphSlots; ++i) repeat |= g_seen_dev[i] == dev + 1 && g_seen[i] == a;
  if (!repeat) {
    g_seen[g_seen_next] = a;
    g_seen_dev[g_seen_next] = dev + 1;
    g_seen_next = (g_seen_next + 1) % kGraphSlots;
  }
  cudaStreamCaptureStatus st = cudaStreamCaptureStatusNone;
  // (a caller whose buffers never settle would pay a capture per call: stop capturing when replays do not follow)
  const bool thrash = g_stats[1] >= 8 && g_stats[0] < 4 * g_stats[1];
  if (!repeat || thrash || (s != nullptr && (cudaStreamIsCapturing(s, &st) != cudaSuccess || st != cudaStreamCaptureStatusNone))) {
    ++g_stats[2];
    return cels_stage_main(a, ws, a.n_rows, s);
  }   // first sighting (or the caller is capturing itself): plain launches
  // captured on a private stream (the caller's may be the legacy default stream, which cannot be captured); the graph
  // itself is launched into the caller's stream
  static cudaStream_t cap[kMaxDevices] = {};
  if (!cap[dev] && cudaStreamCreateWithFlags(&cap[dev], cudaStreamNonBlocking) != cudaSuccess) {
    cap[dev] = nullptr;
    cudaGetLastError();
    return cels_stage_main(a, ws, a.n_rows, s);
  }
  CelsGraph* slot = &g_graphs[0];
  for (auto& g : g_graphs)
    if (!g.exec) { slot = &g; break; } else if (g.used < slot->used) slot = &g;
  if (slot->exec) { cudaGraphExecDestroy(slot->exec); slot->exec = nullptr; }
  cudaGraph_t graph = nullptr;
  if (cudaStreamBeginCapture(cap[dev], cudaStreamCaptureModeThreadLocal) != cudaSuccess) {
    cudaGetLastError();
    return cels_stage_main(a, ws, a.n_rows, s);
  }
  const int rc = cels_stage_main(a, ws, a.n_rows, cap[dev]);
  const cudaError_t e = cudaStreamEndCapture(cap[dev], &graph);
  if (rc != TDSSF_OK || e != cudaSuccess || !graph) {
    if (graph) cudaGraphDestroy(graph);
    cudaGetLastError();
    return rc != TDSSF_OK ? rc : cels_stage_main(a, ws, a.n_rows, s);
  }
  cudaGraphExec_t exec = nullptr;
  const cudaError_t ei = cudaGraphInstantiate(&exec, graph, 0);
  cudaGraphDestroy(graph);
  if (ei != cudaSuccess || !exec) {
    cudaGetLastError();
    return cels_stage_main(a, ws, a.n_rows, s);
  }
  *slot = CelsGraph{a, dev, exec, g_tick};
  ++g_stats[1];
  TDSSF_CUDA(cudaGraphLaunch(exec, s));
  return TDSSF_OK;
}

}  // namespace tdssf

extern "C" int tdssf_cels_fwd_bwd(const float* logits, const int64_t* labels, const uint8_t* mask,
                                  int64_t n_rows, const float* class_w, float lambda_lovasz, int ignore_index,
                                  float* out_scalars, int64_t* n_valid_host, int64_t* preds, int64_t* gt,
                                  float* dlogits, void* workspace, size_t workspace_bytes, void* stream) {
  using namespace tdssf;
  TDSSF_REQUIRE(n_valid_host, "tdssf_cels_fwd_bwd: NULL n_valid_host");
  const CelsArgs a = {logits, labels, mask, n_rows, class_w, lambda_lovasz, ignore_index, out_scalars, preds, gt, dlogits, workspace};
  if (int rc = cels_check(a, workspace_bytes, "tdssf_cels_fwd_bwd")) return rc;
  cudaStream_t s = (cudaStream_t)stream;
  CelsWs ws;
  char* base = reinterpret_cast<char*>(((uintptr_t)workspace + 255) & ~(uintptr_t)255);
  carve(&ws, base, std::max<int64_t>(n_rows, 1));
  if (int rc = cels_stage_count(a, ws, s)) return rc;
  int n_valid = 0;
  *n_valid_host = 0;
  if (n_rows > 0) {
    TDSSF_CUDA(cudaMemcpyAsync(&n_valid, ws.n_valid, sizeof(int), cudaMemcpyDeviceToHost, s));
    TDSSF_CUDA(cudaStreamSynchronize(s));  // the reference's boolean indexing synchronises here as well
    *n_valid_host = n_valid;
  }
  return cels_stage_main(a, ws, n_valid, s);
}

extern "C" void tdssf_cels_graph_stats(int64_t* replays, int64_t* captures, int64_t* plain) {
  std::lock_guard<std::mutex> lock(tdssf::g_graph_mu);
  if (replays) *replays = tdssf::g_stats[0];
  if (captures) *captures = tdssf::g_stats[1];
  if (plain) *plain = tdssf::g_stats[2];
}

extern "C" int tdssf_cels_count(const int64_t* labels, const uint8_t* mask, int64_t n_rows, int ignore_index, int32_t* n_valid_dev,
                                void* workspace, size_t workspace_bytes, void* stream) {
  using namespace tdssf;
  TDSSF_REQUIRE(n_rows >= 0 && n_rows < (int64_t)1 << 27, "tdssf_cels_count: n_rows=%lld out of range", (long long)n_rows);
  TDSSF_REQUIRE(workspace && n_valid_dev && (labels || n_rows == 0), "tdssf_cels_count: NULL pointer");
  TDSSF_REQUIRE(workspace_bytes >= tdssf_cels_workspace_bytes(n_rows), "tdssf_cels_count: workspace too small");
  cudaStream_t s = (cudaStream_t)stream;
  CelsWs ws;
  char* base = reinterpret_cast<char*>(((uintptr_t)workspace + 255) & ~(uintptr_t)255);
  carve(&ws, base, std::max<int64_t>(n_rows, 1));
  CelsArgs a = {};
  a.labels = labels; a.mask = mask; a.n_rows = n_rows; a.ignore = ignore_index;
  if (int rc = cels_stage_count(a, ws, s)) return rc;
  TDSSF_CUDA(cudaMemcpyAsync(n_valid_dev, ws.n_valid, sizeof(int), cudaMemcpyDeviceToDevice, s));
  return TDSSF_OK;
}

extern "C" int tdssf_cels_fwd_bwd_counted(const float* logits, const int64_t* labels, const uint8_t* mask,
                                          int64_t n_rows, const float* class_w, float lambda_lovasz, int ignore_index,
                                          float* out_scalars, int64_t* preds, int64_t* gt,
                                          float* dlogits, void* workspace, size_t workspace_bytes, void* stream) {
  using namespace tdssf;
  const CelsArgs a = {logits, labels, mask, n_rows, class_w, lambda_lovasz, ignore_index, out_scalars, preds, gt, dlogits, workspace};
  if (int rc = cels_check(a, workspace_bytes, "tdssf_cels_fwd_bwd_counted")) return rc;
  CelsWs ws;
  char* base = reinterpret_cast<char*>(((uintptr_t)workspace + 255) & ~(uintptr_t)255);
  carve(&ws, base, std::max<int64_t>(n_rows, 1));
  return cels_stage_main_graphed(a, ws, (cudaStream_t)stream);
}

extern "C" int tdssf_cels_backward(const void* workspace, size_t workspace_bytes, int64_t n_rows, int64_t n_valid,
                                   const float* class_w, const float* coef, float* dlogits, void* stream) {
  using namespace tdssf;
  TDSSF_REQUIRE(workspace && class_w && coef && dlogits, "tdssf_cels_backward: NULL pointer");
  TDSSF_REQUIRE(n_rows >= 0 && n_valid >= 0 && n_valid <= n_rows, "tdssf_cels_backward: bad row counts");
  TDSSF_REQUIRE(workspace_bytes >= tdssf_cels_workspace_bytes(n_rows), "tdssf_cels_backward: workspace too small");
  TDSSF_REQUIRE(((uintptr_t)dlogits & 15) == 0, "tdssf_cels_backward: dlogits must be 16-byte aligned");
  cudaStream_t s = (cudaStream_t)stream;
  if (n_rows == 0) return TDSSF_OK;
  TDSSF_CUDA(cudaMemsetAsync(dlogits, 0, (size_t)n_rows * kC * sizeof(float), s));
  if (n_valid == 0) return TDSSF_OK;
  CelsWs ws;
  char* base = reinterpret_cast<char*>(((uintptr_t)workspace + 255) & ~(uintptr_t)255);
  carve(&ws, base, std::max<int64_t>(n_rows, 1));
  cels_backward_kernel<<<(unsigned)ceil_div(n_valid, 256), 256, 0, s>>>(ws, class_w, 0.f, 0.f, coef, dlogits);
  TDSSF_LAUNCH_CHECK();
  return TDSSF_OK;
}
