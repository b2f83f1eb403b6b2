// f1: voxelizer (utils/lidar.py:83-118, `voxelize_open3d`): the reference ships every cloud to the host, runs Open3D's
// voxel_down_sample and a scikit-learn 5-NN there and ships the voxels back.  Here the whole step stays on the device:
//   1. bounds      per-cloud min corner (fp32 min is exact and order-free)
//   2. insert      voxel index of every point in double, exactly Open3D 0.18's arithmetic
//                  (idx = int(floor((p - (min - size/2)) / size))), into an open-addressing table keyed by the packed
//                  index; the table entry keeps the LOWEST point index of the voxel.  The same pass links every point
//                  into a coarser hash grid for the neighbour search.
//   3. rank        voxels are numbered in the order of their first point (a tile count + scan over "I am the first
//                  point of my voxel" flags): deterministic, unlike the unordered_map order of Open3D.
//   4. sort        stable LSD radix sort of the point indices by voxel number, so that
//   5. centroid    one thread per voxel adds its points in double IN POINT ORDER (the order Open3D's accumulator sees
//                  them) and divides by the count: bit-identical centroids; grid coords = int(floor(centroid / size)).
//   6. 5-NN        a warp per voxel walks the 27 grid cells around the centroid keeping the five nearest points by
//                  (double squared distance, index); accepted when the fifth lies within 0.99 cell.  The points are
//                  linked into kLevels grids whose cells double in size, and a voxel moves to the first coarser grid
//                  that can certify the distance it has found (sparse far-range voxels); what even the coarsest grid
//                  cannot certify is scanned by a whole block.  Intensity = sum_k I[k] * {0.4,0.3,0.2,0.05,0.05}[k]
//                  in double, in that order.
// All of it is integer / hashing / fp64-scalar work bound by latency and atomics, not by HBM bandwidth or tensor cores.
#include <algorithm>
#include "common.cuh"

namespace tdssf {
namespace vox {

constexpr int kTile = 256;                 // points per CTA in the per-point passes
constexpr int kBins = 256;                 // radix 2^8
constexpr int kSortThreads = 256;
constexpr int kKeysPerThread = 8;
constexpr int kSortTile = kSortThreads * kKeysPerThread;
constexpr int kAxisBits = 21;              // packed voxel index: 3 x 21 bits
constexpr unsigned long long kEmpty = ~0ull;
constexpr int kK = 5;
constexpr int kLevels = 6;                 // neighbour grids: cell, 2 cell, ... 32 cell
#ifndef TDSSF_VOX_START
#define TDSSF_VOX_START 48
#endif
// the search starts at the finest grid whose 27 buckets around the centre hold this many points: with n points in the
// 27 c^3 cube about n/6.4 lie within the 0.99 c ball that certifies the result, so 48 makes a second walk rare
constexpr int kStartCount = TDSSF_VOX_START;

struct Ws {
  int* bounds;          // [B][4]  ordered-int encoding of the min corner (init: large)
  int* first;           // [B][T]  lowest point index per table slot (init: large)
  unsigned long long* keys;  // [B][T] packed voxel index (init: empty)
  int* head;            // [kLevels][B][T]  neighbour-grid bucket heads (init: -1)
  int* cnt;             // [kLevels][B][T]  points per bucket (init: 0)
  int* next;            // [kLevels][B][P]  neighbour-grid links
  int* slot;            // [B][P]  table slot of every point
  int* slot_rank;       // [B][T]  voxel number per table slot
  int* tile_cnt;        // [B][tiles]
  int* tile_off;        // [B][tiles]
  uint32_t* key_a; uint32_t* key_b; uint32_t* val_a; uint32_t* val_b;   // [B][P] each
  uint32_t* hist;       // [B][kBins][sort_tiles]
  int* seg;             // [B][P+1] first sorted position of every voxel
  double* cen;          // [B][P][3] centroids in double
  int* todo;            // [B*P] voxels whose 5-NN needs the full scan
  int* counters;        // [0] n_todo, [1] status flags
};

static size_t align_up(size_t x) { return (x + 255) & ~(size_t)255; }
static uint32_t table_size(int P) {
  uint32_t t = 1024;
  while (t < 2u * (uint32_t)P) t <<= 1;
  return t;
}
// returns the bytes needed; fills *ws when base != nullptr.  The first three regions are contiguous per init pattern.
static size_t carve(Ws* ws, char* base, int B, int P) {
  const size_t T = table_size(P), tiles = ceil_div(P, kTile), stiles = ceil_div(P, kSortTile);
  size_t off = 0;
  auto take = [&](size_t bytes) { char* p = base ? base + off : nullptr; off += align_up(bytes); return p; };
  char* p;
  p = take((size_t)B * 4 * 4); if (ws) ws->bounds = (int*)p;                 // 0x7f region
  p = take((size_t)B * T * 4); if (ws) ws->first = (int*)p;
  p = take((size_t)B * T * 8); if (ws) ws->keys = (unsigned long long*)p;    // 0xff region
  p = take((size_t)kLevels * B * T * 4); if (ws) ws->head = (int*)p;
  p = take(256); if (ws) ws->counters = (int*)p;                             // 0x00 region
  p = take((size_t)kLevels * B * T * 4); if (ws) ws->cnt = (int*)p;
  p = take((size_t)kLevels * B * P * 4); if (ws) ws->next = (int*)p;
  p = take((size_t)B * P * 4); if (ws) ws->slot = (int*)p;
  p = take((size_t)B * T * 4); if (ws) ws->slot_rank = (int*)p;
  p = take((size_t)B * tiles * 4); if (ws) ws->tile_cnt = (int*)p;
  p = take((size_t)B * tiles * 4); if (ws) ws->tile_off = (int*)p;
  p = take((size_t)B * P * 4); if (ws) ws->key_a = (uint32_t*)p;
  p = take((size_t)B * P * 4); if (ws) ws->key_b = (uint32_t*)p;
  p = take((size_t)B * P * 4); if (ws) ws->val_a = (uint32_t*)p;
  p = take((size_t)B * P * 4); if (ws) ws->val_b = (uint32_t*)p;
  p = take((size_t)B * kBins * stiles * 4); if (ws) ws->hist = (uint32_t*)p;
  p = take((size_t)B * (P + 1) * 4); if (ws) ws->seg = (int*)p;
  p = take((size_t)B * P * 3 * 8); if (ws) ws->cen = (double*)p;
  p = take((size_t)B * P * 4); if (ws) ws->todo = (int*)p;
  return off + 256;
}

// monotone map float -> int (signed compare), and back
__device__ __forceinline__ int enc_ordered(float f) {
  const int b = __float_as_int(f);
  return b >= 0 ? b : b ^ 0x7fffffff;
}
__device__ __forceinline__ float dec_ordered(int e) { return __int_as_float(e >= 0 ? e : e ^ 0x7fffffff); }

__device__ __forceinline__ uint32_t mix64(unsigned long long k) {
  k ^= k >> 33; k *= 0xff51afd7ed558ccdull; k ^= k >> 33; k *= 0xc4ceb9fe1a85ec53ull; k ^= k >> 33;
  return (uint32_t)k;
}
// neighbour-grid cell of a coordinate: the same double expression for points and queries (monotone in x)
__device__ __forceinline__ int cell_of(double x, double inv_cell) { return (int)floor(x * inv_cell); }
__device__ __forceinline__ uint32_t cell_hash(int cx, int cy, int cz) {
  return (uint32_t)cx * 73856093u ^ (uint32_t)cy * 19349663u ^ (uint32_t)cz * 83492791u;
}

// ---- 1. bounds ----------------------------------------------------------------------------------
__global__ void __launch_bounds__(kTile)
bounds_kernel(const float* __restrict__ pts, int64_t stride, int P, int* __restrict__ bounds) {
  const int b = blockIdx.y, i = blockIdx.x * kTile + threadIdx.x;
  int e[3] = {0x7f7f7f7f, 0x7f7f7f7f, 0x7f7f7f7f};
  if (i < P) {
    const float* q = pts + ((int64_t)b * P + i) * stride;
#pragma unroll
    for (int a = 0; a < 3; ++a) e[a] = enc_ordered(__ldg(q + a));
  }
#pragma unroll
  for (int a = 0; a < 3; ++a) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) e[a] = min(e[a], __shfl_xor_sync(0xffffffffu, e[a], o));
  }
  if ((threadIdx.x & 31) == 0) {
#pragma unroll
    for (int a = 0; a < 3; ++a) atomicMin(bounds + b * 4 + a, e[a]);
  }
}

// ---- 2. voxel table + neighbour grid ------------------------------------------------------------
__global__ void __launch_bounds__(kTile)
insert_kernel(const float* __restrict__ pts, int64_t stride, int P, int B, double voxel_size, double inv_cell, uint32_t mask, Ws ws) {
  const int b = blockIdx.y, i = blockIdx.x * kTile + threadIdx.x;
  if (i >= P) return;
  const float* q = pts + ((int64_t)b * P + i) * stride;
  const double x = (double)__ldg(q), y = (double)__ldg(q + 1), z = (double)__ldg(q + 2);
  const double p[3] = {x, y, z};
  unsigned long long key = 0;
  bool bad = false;
#pragma unroll
  for (int a = 0; a < 3; ++a) {
    // Open3D: voxel_min_bound = min_bound - voxel_size * 0.5; ref = (p - voxel_min_bound) / voxel_size; int(floor(ref))
    const double lo = __dsub_rn((double)dec_ordered(ws.bounds[b * 4 + a]), __dmul_rn(voxel_size, 0.5));
    const double ref = __ddiv_rn(__dsub_rn(p[a], lo), voxel_size);
    const double f = floor(ref);
    long long idx = (f >= 0.0 && f < 2097152.0) ? (long long)f : -1;
    if (idx < 0) { bad = true; idx = 0; }   // NaN / inf coordinate or more than 2^21 voxels along an axis
    key |= (unsigned long long)idx << (a * kAxisBits);
  }
  if (bad) atomicOr(ws.counters + 1, 1);
  const size_t T = (size_t)mask + 1;
  unsigned long long* keys = ws.keys + (size_t)b * T;
  uint32_t s = mix64(key) & mask;
  while (true) {
    const unsigned long long old = atomicCAS(keys + s, kEmpty, key);
    if (old == kEmpty || old == key) break;
    s = (s + 1) & mask;
  }
  atomicMin(ws.first + (size_t)b * T + s, i);
  ws.slot[(size_t)b * P + i] = (int)s;
  // neighbour grids over the points
  double inv = inv_cell;
#pragma unroll
  for (int l = 0; l < kLevels; ++l, inv *= 0.5) {
    const uint32_t h = cell_hash(cell_of(x, inv), cell_of(y, inv), cell_of(z, inv)) & mask;
    ws.next[((size_t)l * B + b) * P + i] = atomicExch(ws.head + ((size_t)l * B + b) * T + h, i);
    atomicAdd(ws.cnt + ((size_t)l * B + b) * T + h, 1);
  }
}

// ---- 3. voxel numbers in order of first appearance ------------------------------------------------
__global__ void __launch_bounds__(kTile)
flag_count_kernel(int P, uint32_t mask, Ws ws, int tiles) {
  const int b = blockIdx.y, i = blockIdx.x * kTile + threadIdx.x;
  const size_t T = (size_t)mask + 1;
  const int f = i < P && ws.first[(size_t)b * T + ws.slot[(size_t)b * P + i]] == i;
  const int c = __syncthreads_count(f);
  if (threadIdx.x == 0) ws.tile_cnt[(size_t)b * tiles + blockIdx.x] = c;
}

__device__ __forceinline__ int warp_incl_scan(int v) {
  const int lane = threadIdx.x & 31;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    const int t = __shfl_up_sync(0xffffffffu, v, o);
    if (lane >= o) v += t;
  }
  return v;
}
__device__ __forceinline__ int block_excl_scan256(int v, int* s_warp /*[9]*/, int* total) {
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  const int incl = warp_incl_scan(v);
  if (lane == 31) s_warp[w] = incl;
  __syncthreads();
  if (w == 0) {
    const int x = lane < 8 ? s_warp[lane] : 0;
    const int xi = warp_incl_scan(x);
    if (lane < 8) s_warp[lane] = xi - x;
    if (lane == 7) s_warp[8] = xi;
  }
  __syncthreads();
  const int r = s_warp[w] + incl - v;
  *total = s_warp[8];
  __syncthreads();
  return r;
}

// one CTA per cloud: exclusive scan of the tile counts; the total is the cloud's voxel count
__global__ void __launch_bounds__(256)
scan_tiles_kernel(Ws ws, int tiles, int32_t* __restrict__ n_voxels) {
  __shared__ int s_warp[9];
  const int b = blockIdx.x;
  const int* in = ws.tile_cnt + (size_t)b * tiles;
  int* out = ws.tile_off + (size_t)b * tiles;
  int carry = 0;
  for (int base = 0; base < tiles; base += 256) {
    const int i = base + threadIdx.x;
    const int v = i < tiles ? in[i] : 0;
    int tot;
    const int e = block_excl_scan256(v, s_warp, &tot);
    if (i < tiles) out[i] = carry + e;
    carry += tot;
  }
  if (threadIdx.x == 0) n_voxels[b] = carry;
}

__global__ void __launch_bounds__(kTile)
rank_kernel(int P, uint32_t mask, Ws ws, int tiles) {
  __shared__ int s_warp[9];
  const int b = blockIdx.y, i = blockIdx.x * kTile + threadIdx.x;
  const size_t T = (size_t)mask + 1;
  const int s = i < P ? ws.slot[(size_t)b * P + i] : 0;
  const int f = i < P && ws.first[(size_t)b * T + s] == i;
  int tot;
  const int e = block_excl_scan256(f, s_warp, &tot);
  if (f) ws.slot_rank[(size_t)b * T + s] = ws.tile_off[(size_t)b * tiles + blockIdx.x] + e;
}

__global__ void __launch_bounds__(kTile)
keys_kernel(int P, uint32_t mask, Ws ws) {
  const int b = blockIdx.y, i = blockIdx.x * kTile + threadIdx.x;
  if (i >= P) return;
  const size_t T = (size_t)mask + 1;
  ws.key_a[(size_t)b * P + i] = (uint32_t)ws.slot_rank[(size_t)b * T + ws.slot[(size_t)b * P + i]];
  ws.val_a[(size_t)b * P + i] = (uint32_t)i;
}

// ---- 4. stable LSD radix sort of (voxel number, point index), one segment per cloud ------------------
// hist layout [cloud][bin][tile]: one exclusive scan per cloud gives the global offset of every (bin, tile) pair
__global__ void __launch_bounds__(kSortThreads)
sort_hist_kernel(const uint32_t* __restrict__ keys, int N, int shift, int tiles, uint32_t* __restrict__ hist) {
  const int seg = blockIdx.y, tile = blockIdx.x;
  __shared__ uint32_t s_h[kBins];
  s_h[threadIdx.x] = 0;
  __syncthreads();
  const uint32_t* k = keys + (int64_t)seg * N;
  const int base = tile * kSortTile;
#pragma unroll
  for (int j = 0; j < kKeysPerThread; ++j) {
    const int idx = base + j * kSortThreads + threadIdx.x;
    if (idx < N) atomicAdd(&s_h[(k[idx] >> shift) & (kBins - 1)], 1u);
  }
  __syncthreads();
  hist[((int64_t)seg * kBins + threadIdx.x) * tiles + tile] = s_h[threadIdx.x];
}

__global__ void __launch_bounds__(256)
sort_scan_kernel(uint32_t* __restrict__ hist, int tiles) {
  __shared__ int s_warp[9];
  uint32_t* h = hist + (int64_t)blockIdx.x * kBins * tiles;
  const int64_t n = (int64_t)kBins * tiles;
  int carry = 0;
  for (int64_t base = 0; base < n; base += 2048) {
    const int64_t i0 = base + threadIdx.x * 8;
    uint32_t v[8];
    int sum = 0;
#pragma unroll
    for (int j = 0; j < 8; ++j) { v[j] = (i0 + j < n) ? h[i0 + j] : 0u; sum += (int)v[j]; }
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

// warp w owns 256 consecutive keys of the tile, taken as 8 rounds of 32 consecutive keys, so (round, lane) order is key
// order and the ranking below is stable
__global__ void __launch_bounds__(kSortThreads)
sort_scatter_kernel(const uint32_t* __restrict__ keys_in, const uint32_t* __restrict__ vals_in, uint32_t* __restrict__ keys_out,
                    uint32_t* __restrict__ vals_out, int N, int shift, int tiles, const uint32_t* __restrict__ hist) {
  const int seg = blockIdx.y, tile = blockIdx.x;
  constexpr int kWarps = kSortThreads / 32;
  __shared__ uint32_t s_cnt[kWarps][kBins];
  __shared__ uint32_t s_base[kBins];
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  for (int i = threadIdx.x; i < kWarps * kBins; i += kSortThreads) (&s_cnt[0][0])[i] = 0;
  s_base[threadIdx.x] = hist[((int64_t)seg * kBins + threadIdx.x) * tiles + tile];
  __syncthreads();
  const int64_t so = (int64_t)seg * N;
  const int wbase = tile * kSortTile + w * (32 * kKeysPerThread);
  uint32_t key[kKeysPerThread], val[kKeysPerThread], rank[kKeysPerThread];
#pragma unroll
  for (int j = 0; j < kKeysPerThread; ++j) {
    const int idx = wbase + j * 32 + lane;
    const bool in = idx < N;
    key[j] = in ? keys_in[so + idx] : 0xffffffffu;
    val[j] = in ? vals_in[so + idx] : 0u;
    const uint32_t d = (key[j] >> shift) & (kBins - 1);
    const uint32_t act = __ballot_sync(0xffffffffu, in);
    uint32_t r = 0;
    if (in) {
      const uint32_t peers = __match_any_sync(act, d);
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
  {
    uint32_t run = 0;
#pragma unroll
    for (int ww = 0; ww < kWarps; ++ww) {
      const uint32_t c = s_cnt[ww][threadIdx.x];
      s_cnt[ww][threadIdx.x] = run;
      run += c;
    }
  }
  __syncthreads();
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

// first sorted position of every voxel (and the end marker of the last one)
__global__ void __launch_bounds__(kTile)
seg_start_kernel(const uint32_t* __restrict__ keys, int P, int* __restrict__ seg, const int32_t* __restrict__ n_voxels) {
  const int b = blockIdx.y, j = blockIdx.x * kTile + threadIdx.x;
  if (j >= P) return;
  const uint32_t* k = keys + (size_t)b * P;
  int* s = seg + (size_t)b * (P + 1);
  if (j == 0 || k[j] != k[j - 1]) s[k[j]] = j;
  if (j == P - 1) s[n_voxels[b]] = P;
}

// ---- 5. centroids ---------------------------------------------------------------------------------
__global__ void __launch_bounds__(kTile)
centroid_kernel(const float* __restrict__ pts, int64_t stride, int P, double voxel_size, const uint32_t* __restrict__ vals,
                Ws ws, const int32_t* __restrict__ n_voxels, float* __restrict__ raw, int32_t* __restrict__ coords) {
  const int b = blockIdx.y, v = blockIdx.x * kTile + threadIdx.x;
  if (v >= P) return;
  float* r = raw + ((size_t)b * P + v) * 4;
  int32_t* c = coords + ((size_t)b * P + v) * 3;
  if (v >= n_voxels[b]) {  // padding rows, like the zero padding of LiDARFeatureEncoder.forward (utils/lidar.py:67-70)
    r[0] = r[1] = r[2] = r[3] = 0.f;
    c[0] = c[1] = c[2] = 0;
    return;
  }
  const int* s = ws.seg + (size_t)b * (P + 1);
  const int j0 = s[v], j1 = s[v + 1];
  const uint32_t* order = vals + (size_t)b * P;
  double sx = 0.0, sy = 0.0, sz = 0.0;
  for (int j = j0; j < j1; ++j) {  // ascending point index (stable sort): Open3D's accumulation order
    const float* q = pts + ((int64_t)b * P + order[j]) * stride;
    sx = __dadd_rn(sx, (double)__ldg(q));
    sy = __dadd_rn(sy, (double)__ldg(q + 1));
    sz = __dadd_rn(sz, (double)__ldg(q + 2));
  }
  const double n = (double)(j1 - j0);
  const double cx = __ddiv_rn(sx, n), cy = __ddiv_rn(sy, n), cz = __ddiv_rn(sz, n);
  double* cen = ws.cen + ((size_t)b * P + v) * 3;
  cen[0] = cx; cen[1] = cy; cen[2] = cz;
  r[0] = (float)cx; r[1] = (float)cy; r[2] = (float)cz;
  r[3] = 0.f;
  c[0] = (int32_t)floor(__ddiv_rn(cx, voxel_size));   // np.floor(down_points / voxel_size).astype(np.int32)
  c[1] = (int32_t)floor(__ddiv_rn(cy, voxel_size));
  c[2] = (int32_t)floor(__ddiv_rn(cz, voxel_size));
}

// ---- 6. five nearest points per voxel centre --------------------------------------------------------
struct Top5 {
  double d[kK];
  int i[kK];
  __device__ __forceinline__ void init() {
#pragma unroll
    for (int k = 0; k < kK; ++k) { d[k] = INFINITY; i[k] = 0x7fffffff; }
  }
  // insert keeping (distance, index) ascending
  __device__ __forceinline__ void push(double nd, int ni) {
    if (!(nd < d[kK - 1] || (nd == d[kK - 1] && ni < i[kK - 1]))) return;
    d[kK - 1] = nd; i[kK - 1] = ni;
#pragma unroll
    for (int k = kK - 1; k > 0; --k) {
      const bool sw = d[k] < d[k - 1] || (d[k] == d[k - 1] && i[k] < i[k - 1]);
      if (sw) {
        const double td = d[k]; d[k] = d[k - 1]; d[k - 1] = td;
        const int ti = i[k]; i[k] = i[k - 1]; i[k - 1] = ti;
      }
    }
  }
  __device__ __forceinline__ void pop() {
#pragma unroll
    for (int k = 0; k < kK - 1; ++k) { d[k] = d[k + 1]; i[k] = i[k + 1]; }
    d[kK - 1] = INFINITY; i[kK - 1] = 0x7fffffff;
  }
};

__device__ __forceinline__ double dist2(double qx, double qy, double qz, const float* p) {
  const double dx = __dsub_rn(qx, (double)__ldg(p)), dy = __dsub_rn(qy, (double)__ldg(p + 1)), dz = __dsub_rn(qz, (double)__ldg(p + 2));
  return __dadd_rn(__dadd_rn(__dmul_rn(dx, dx), __dmul_rn(dy, dy)), __dmul_rn(dz, dz));
}

// the warp's five smallest (distance, index) pairs out of the lanes' lists, nearest first, to every lane
__device__ __forceinline__ void warp_merge(Top5& t, double* od, int* oi) {
#pragma unroll
  for (int k = 0; k < kK; ++k) {
    double bd = t.d[0];
    int bi = t.i[0];
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      const double xd = __shfl_xor_sync(0xffffffffu, bd, o);
      const int xi = __shfl_xor_sync(0xffffffffu, bi, o);
      if (xd < bd || (xd == bd && xi < bi)) { bd = xd; bi = xi; }
    }
    od[k] = bd; oi[k] = bi;
    if (t.i[0] == bi && t.d[0] == bd) t.pop();   // a point sits in exactly one lane's list
  }
}

// sum_k I[k] * w[k], double, left to right ((pts[idx,3] * weights).sum(axis=1), utils/lidar.py:101-102)
__device__ __forceinline__ float weighted_intensity(const float* __restrict__ cloud, int64_t stride, int P, const int* nbr) {
  const double w[kK] = {0.4, 0.3, 0.2, 0.05, 0.05};
  double acc = 0.0;
#pragma unroll
  for (int k = 0; k < kK; ++k) {
    // (no neighbour: only for a centre with a non-finite coordinate, which the status word reports)
    const double v = (unsigned)nbr[k] < (unsigned)P ? (double)__ldg(cloud + (int64_t)nbr[k] * stride + 3) : (double)NAN;
    acc = __dadd_rn(acc, __dmul_rn(v, w[k]));
  }
  return (float)acc;
}

constexpr int kQueryThreads = 256;
__global__ void __launch_bounds__(kQueryThreads)
knn_grid_kernel(const float* __restrict__ pts, int64_t stride, int P, int B, double cell, uint32_t mask, Ws ws,
                const int32_t* __restrict__ n_voxels, float* __restrict__ raw) {
  const int b = blockIdx.y, lane = threadIdx.x & 31;
  const int nv = n_voxels[b];
  const size_t T = (size_t)mask + 1;
  const float* cloud = pts + (int64_t)b * P * stride;
  for (int v = blockIdx.x * (kQueryThreads / 32) + (threadIdx.x >> 5); v < nv; v += gridDim.x * (kQueryThreads / 32)) {
    const double* cen = ws.cen + ((size_t)b * P + v) * 3;
    const double qx = cen[0], qy = cen[1], qz = cen[2];
    double d[kK];
    int nb[kK];
    bool done = false;
    // bucket counts of all grids: one round of independent loads instead of walking the lists of grids that are too fine
    // (or too coarse) for this neighbourhood
    int l = kLevels - 1;
    {
      int c[kLevels];
#pragma unroll
      for (int k = 0; k < kLevels; ++k) {
        const double inv = (1.0 / cell) / (double)(1 << k);
        const int tx = cell_of(qx, inv) + lane % 3 - 1, ty = cell_of(qy, inv) + (lane / 3) % 3 - 1, tz = cell_of(qz, inv) + lane / 9 - 1;
        c[k] = lane < 27 ? ws.cnt[((size_t)k * B + b) * T + (cell_hash(tx, ty, tz) & mask)] : 0;
      }
#pragma unroll
      for (int k = kLevels - 1; k >= 0; --k) {
        if (__reduce_add_sync(0xffffffffu, c[k]) >= kStartCount) l = k;
      }
    }
    while (l < kLevels) {
      const double c = cell * (double)(1 << l), inv = (1.0 / cell) / (double)(1 << l);   // powers of two: exact, as in insert_kernel
      const int* head = ws.head + ((size_t)l * B + b) * T;
      const int* next = ws.next + ((size_t)l * B + b) * P;
      const int cx = cell_of(qx, inv), cy = cell_of(qy, inv), cz = cell_of(qz, inv);
      Top5 t;
      t.init();
      if (lane < 27) {  // one of the 27 cells around the centre per lane
        const int tx = cx + lane % 3 - 1, ty = cy + (lane / 3) % 3 - 1, tz = cz + lane / 9 - 1;
        for (int n = head[cell_hash(tx, ty, tz) & mask]; n >= 0; n = next[n]) {
          const float* p = cloud + (int64_t)n * stride;
          // buckets are shared by colliding cells: only the points of this lane's cell (each point has exactly one)
          if (cell_of((double)__ldg(p), inv) != tx || cell_of((double)__ldg(p + 1), inv) != ty ||
              cell_of((double)__ldg(p + 2), inv) != tz)
            continue;
          t.push(dist2(qx, qy, qz, p), n);
        }
      }
      warp_merge(t, d, nb);
      const double accept = 0.99 * c;
      if (d[kK - 1] <= accept * accept) { done = true; break; }   // everything nearer than 0.99 cell lies in the 27 cells
      // the fifth neighbour is at most sqrt(d[4]) away (if five were found): go to the first grid that can certify that
      ++l;
      while (l < kLevels && d[kK - 1] < INFINITY && d[kK - 1] > 0.9801 * cell * cell * (double)(1 << l) * (double)(1 << l)) ++l;
    }
    if (lane == 0) {
      if (done) raw[((size_t)b * P + v) * 4 + 3] = weighted_intensity(cloud, stride, P, nb);
      else ws.todo[atomicAdd(ws.counters, 1)] = b * P + v;
    }
  }
}

// a block per unresolved voxel: all P points
constexpr int kScanThreads = 256;
__global__ void __launch_bounds__(kScanThreads)
knn_scan_kernel(const float* __restrict__ pts, int64_t stride, int P, Ws ws, float* __restrict__ raw) {
  __shared__ double s_d[kScanThreads / 32][kK];
  __shared__ int s_i[kScanThreads / 32][kK];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int n = ws.counters[0];
  for (int w = blockIdx.x; w < n; w += gridDim.x) {
    const int gv = ws.todo[w], b = gv / P;
    const double* cen = ws.cen + (size_t)gv * 3;
    const double qx = cen[0], qy = cen[1], qz = cen[2];
    const float* cloud = pts + (int64_t)b * P * stride;
    Top5 t;
    t.init();
    for (int i = threadIdx.x; i < P; i += kScanThreads) t.push(dist2(qx, qy, qz, cloud + (int64_t)i * stride), i);
    double d[kK];
    int nb[kK];
    warp_merge(t, d, nb);
    if (lane == 0) {
#pragma unroll
      for (int k = 0; k < kK; ++k) { s_d[warp][k] = d[k]; s_i[warp][k] = nb[k]; }
    }
    __syncthreads();
    if (warp == 0) {
      t.init();
      if (lane < kScanThreads / 32) {
#pragma unroll
        for (int k = 0; k < kK; ++k) { t.d[k] = s_d[lane][k]; t.i[k] = s_i[lane][k]; }   // already ascending
      }
      warp_merge(t, d, nb);
      if (lane == 0) raw[(size_t)gv * 4 + 3] = weighted_intensity(cloud, stride, P, nb);
    }
    __syncthreads();
  }
}

}  // namespace vox
}  // namespace tdssf

extern "C" size_t tdssf_voxelize_workspace_bytes(int B, int P) {
  if (B <= 0 || P <= 0) return 256;
  return tdssf::vox::carve(nullptr, nullptr, B, P);
}

extern "C" int tdssf_voxelize(const float* pts, int64_t point_stride_floats, int B, int P, int has_intensity, double voxel_size,
                              double knn_cell, float* voxel_raw, int32_t* voxel_coords, int32_t* n_voxels, int32_t* status,
                              void* workspace, size_t workspace_bytes, void* stream) {
  using namespace tdssf;
  using namespace tdssf::vox;
  TDSSF_REQUIRE(pts && voxel_raw && voxel_coords && n_voxels, "tdssf_voxelize: NULL pointer");
  TDSSF_REQUIRE(B >= 0 && P >= 0, "tdssf_voxelize: negative size");
  TDSSF_REQUIRE(point_stride_floats >= (has_intensity ? 4 : 3), "tdssf_voxelize: point stride of %lld floats is too short",
                (long long)point_stride_floats);
  TDSSF_REQUIRE(voxel_size > 0.0 && voxel_size < 1e300, "tdssf_voxelize: voxel_size must be positive (Open3D: 'voxel_size <= 0')");
  if (B == 0 || P == 0) return TDSSF_OK;
  TDSSF_REQUIRE(!has_intensity || P >= kK, "tdssf_voxelize: %d neighbours asked of %d points (scikit-learn raises ValueError)", kK, P);
  TDSSF_REQUIRE((int64_t)B * P < (1ll << 30), "tdssf_voxelize: B*P exceeds the index range");
  if (!(knn_cell > 0.0)) knn_cell = std::max(2.0 * voxel_size, 0.25);
  const size_t need = carve(nullptr, nullptr, B, P);
  TDSSF_REQUIRE(workspace && ((uintptr_t)workspace & 255) == 0 && workspace_bytes >= need,
                "tdssf_voxelize: 256-byte aligned workspace of %zu bytes needed (tdssf_voxelize_workspace_bytes), got %zu", need,
                workspace_bytes);
  cudaStream_t s = (cudaStream_t)stream;
  Ws ws;
  carve(&ws, reinterpret_cast<char*>(workspace), B, P);
  const uint32_t T = table_size(P), mask = T - 1;
  const int tiles = (int)ceil_div(P, kTile), stiles = (int)ceil_div(P, kSortTile);
  TDSSF_CUDA(cudaMemsetAsync(ws.bounds, 0x7f, (char*)ws.keys - (char*)ws.bounds, s));
  TDSSF_CUDA(cudaMemsetAsync(ws.keys, 0xff, (char*)ws.counters - (char*)ws.keys, s));
  TDSSF_CUDA(cudaMemsetAsync(ws.counters, 0, (char*)ws.next - (char*)ws.counters, s));
  const dim3 pgrid((unsigned)tiles, (unsigned)B);
  bounds_kernel<<<pgrid, kTile, 0, s>>>(pts, point_stride_floats, P, ws.bounds);
  TDSSF_LAUNCH_CHECK();
  insert_kernel<<<pgrid, kTile, 0, s>>>(pts, point_stride_floats, P, B, voxel_size, 1.0 / knn_cell, mask, ws);
  TDSSF_LAUNCH_CHECK();
  flag_count_kernel<<<pgrid, kTile, 0, s>>>(P, mask, ws, tiles);
  TDSSF_LAUNCH_CHECK();
  scan_tiles_kernel<<<(unsigned)B, 256, 0, s>>>(ws, tiles, n_voxels);
  TDSSF_LAUNCH_CHECK();
  rank_kernel<<<pgrid, kTile, 0, s>>>(P, mask, ws, tiles);
  TDSSF_LAUNCH_CHECK();
  keys_kernel<<<pgrid, kTile, 0, s>>>(P, mask, ws);
  TDSSF_LAUNCH_CHECK();
  int bits = 1;
  while ((1ll << bits) < P) ++bits;   // voxel numbers are < P
  uint32_t *ka = ws.key_a, *kb = ws.key_b, *va = ws.val_a, *vb = ws.val_b;
  for (int shift = 0; shift < bits; shift += 8) {
    sort_hist_kernel<<<dim3((unsigned)stiles, (unsigned)B), kSortThreads, 0, s>>>(ka, P, shift, stiles, ws.hist);
    TDSSF_LAUNCH_CHECK();
    sort_scan_kernel<<<(unsigned)B, 256, 0, s>>>(ws.hist, stiles);
    TDSSF_LAUNCH_CHECK();
    sort_scatter_kernel<<<dim3((unsigned)stiles, (unsigned)B), kSortThreads, 0, s>>>(ka, va, kb, vb, P, shift, stiles, ws.hist);
    TDSSF_LAUNCH_CHECK();
    std::swap(ka, kb);
    std::swap(va, vb);
  }
  seg_start_kernel<<<pgrid, kTile, 0, s>>>(ka, P, ws.seg, n_voxels);
  TDSSF_LAUNCH_CHECK();
  centroid_kernel<<<pgrid, kTile, 0, s>>>(pts, point_stride_floats, P, voxel_size, va, ws, n_voxels, voxel_raw, voxel_coords);
  TDSSF_LAUNCH_CHECK();
  if (has_intensity) {
    const unsigned qblocks = (unsigned)std::min<int64_t>(ceil_div((int64_t)P * 32, kQueryThreads), (int64_t)num_sms() * 16);
    knn_grid_kernel<<<dim3(qblocks, (unsigned)B), kQueryThreads, 0, s>>>(pts, point_stride_floats, P, B, knn_cell, mask, ws, n_voxels,
                                                                        voxel_raw);
    TDSSF_LAUNCH_CHECK();
    knn_scan_kernel<<<(unsigned)num_sms() * 8, kScanThreads, 0, s>>>(pts, point_stride_floats, P, ws, voxel_raw);
    TDSSF_LAUNCH_CHECK();
  }
  if (status) TDSSF_CUDA(cudaMemcpyAsync(status, ws.counters + 1, 4, cudaMemcpyDeviceToDevice, s));
  return TDSSF_OK;
}
