// One-shot sum all-reduce over NVLink peer memory for the tensor-parallel decode step.
//
// The reference has no tensor parallelism (vllm_metal/platform.py:524-529 pins world size 1);
// SURVEY.md 8(e) shards this path over heads / MLP columns, which leaves one [M, hidden] sum after
// o_proj and one after down_proj per layer -- 512 KiB at the headline shape, far below the size
// where a ring pays, and latency-bound in NCCL (~20 us each, 64 per step).  This kernel does the
// exchange itself through CUDA-IPC mapped peer buffers (NVSwitch gives every pair full bandwidth):
//
//   CTA b owns slice b of the vector on EVERY rank.
//   1. copy my slice of x into my half of the symmetric staging buffer   (local HBM)
//   2. release-store the call's epoch into flags[b][rank] of every peer  (NVLink, 4 B each)
//   3. acquire-spin until every peer's epoch shows up in my flags[b][*]
//   4. read slice b from every peer's staging buffer (16-byte NVLink loads), add in RANK ORDER in
//      fp32 (bit-identical result on every rank), write x in place.
//
// Two staging halves alternate by epoch parity, so no trailing barrier is needed: a peer can only
// overwrite half h two calls later, and it cannot pass the barrier of the call in between before
// I have signalled it, i.e. before I finished reading.  The epoch lives in device memory (one
// counter per CTA), so the kernel replays inside a CUDA graph unchanged.
#include <cstring>

#include "common.cuh"
#include "../../include/b200_decoder_ops.h"

namespace b200 {
namespace ar {

constexpr int kMaxWorld = 8;
constexpr int kThreads = 512;

struct Params {
  void* x;                    // [numel] in place
  char* symm[kMaxWorld];      // staging buffers of every rank (peer-mapped), 2 * half_bytes each
  uint32_t* flags[kMaxWorld];  // [grid][world] epoch slots of every rank (peer-mapped)
  uint32_t* epoch;            // local [grid]
  int64_t nvec;               // 16-byte vectors in x
  int64_t half_bytes;
  int rank, world;
};

__device__ __forceinline__ void st_release_sys(uint32_t* p, uint32_t v) {
  asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ uint32_t ld_acquire_sys(const uint32_t* p) {
  uint32_t v;
  asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
// peer data: a COHERENT 128-bit load (relaxed, system scope).  The staging halves are rewritten by
// the peers every second call while kernels of this rank may still be resident (PDL), so the
// non-coherent .nc path -- which presumes the data is read-only for the kernel's lifetime -- is not
// allowed here; the acquire + __syncthreads below order these loads after the peer's release.
__device__ __forceinline__ uint4 ld_peer_u4(const void* p) {
  uint4 r;
  asm volatile("ld.relaxed.sys.global.v4.u32 {%0,%1,%2,%3}, [%4];"
               : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w)
               : "l"(p)
               : "memory");
  return r;
}

template <typename T>
__device__ __forceinline__ void accumulate(float (&acc)[16 / sizeof(T)], const uint4& v) {
  const T* e = reinterpret_cast<const T*>(&v);
#pragma unroll
  for (int i = 0; i < (int)(16 / sizeof(T)); ++i) acc[i] += to_float(e[i]);
}

template <typename T, int WORLD>
__global__ void __launch_bounds__(kThreads) allreduce_oneshot_kernel(const Params p) {
  pdl_trigger();
  pdl_wait();
  constexpr int VE = 16 / (int)sizeof(T);
  const int b = blockIdx.x;
  const uint32_t e = p.epoch[b] + 1u;
  // slices are whole 128-byte lines (8 vectors): no two CTAs ever share a line of a peer's staging buffer
  const int64_t per = ((p.nvec + gridDim.x - 1) / gridDim.x + 7) / 8 * 8;
  const int64_t lo = b * per;
  const int64_t hi = lo + per < p.nvec ? lo + per : p.nvec;  // (lo may exceed nvec: empty slice, flags still exchanged)
  const int64_t half_off = (int64_t)(e & 1u) * p.half_bytes;
  uint4* xv = reinterpret_cast<uint4*>(p.x);
  uint4* mine = reinterpret_cast<uint4*>(p.symm[p.rank] + half_off);

  // 1. publish my slice.  No per-thread fence: bar.sync orders every thread's stores before the
  // signalling threads' st.release.sys, which is cumulative over them (PTX memory model).
  for (int64_t v = lo + threadIdx.x; v < hi; v += kThreads) mine[v] = xv[v];
  __syncthreads();
  // 2. + 3. per-CTA flag exchange with every peer
  if (threadIdx.x < WORLD && (int)threadIdx.x != p.rank) {
    const int peer = threadIdx.x;
    st_release_sys(p.flags[peer] + (size_t)b * WORLD + p.rank, e);
    const uint32_t* slot = p.flags[p.rank] + (size_t)b * WORLD + peer;
    uint32_t spins = 0;
    while ((int32_t)(ld_acquire_sys(slot) - e) < 0) {
      if (++spins > (1u << 27)) __trap();  // a dead peer must fault this rank, not hang the box
    }
  }
  __syncthreads();
  // 4. gather + sum in rank order
  for (int64_t v = lo + threadIdx.x; v < hi; v += kThreads) {
    uint4 part[WORLD];
#pragma unroll
    for (int r = 0; r < WORLD; ++r)
      part[r] = (r == p.rank) ? xv[v] : ld_peer_u4(reinterpret_cast<const uint4*>(p.symm[r] + half_off) + v);
    float acc[VE];
#pragma unroll
    for (int i = 0; i < VE; ++i) acc[i] = 0.f;
#pragma unroll
    for (int r = 0; r < WORLD; ++r) accumulate<T>(acc, part[r]);
    uint4 o;
    T* oe = reinterpret_cast<T*>(&o);
#pragma unroll
    for (int i = 0; i < VE; ++i) oe[i] = from_float<T>(acc[i]);
    xv[v] = o;
  }
  __syncthreads();
  if (threadIdx.x == 0) p.epoch[b] = e;
}

template <typename T>
static int launch(const Params& p, int grid, cudaStream_t st) {
  switch (p.world) {
    case 2: B200_CUDA_OK(launch_pdl(allreduce_oneshot_kernel<T, 2>, dim3(grid), dim3(kThreads), 0, st, p)); break;
    case 4: B200_CUDA_OK(launch_pdl(allreduce_oneshot_kernel<T, 4>, dim3(grid), dim3(kThreads), 0, st, p)); break;
    case 8: B200_CUDA_OK(launch_pdl(allreduce_oneshot_kernel<T, 8>, dim3(grid), dim3(kThreads), 0, st, p)); break;
    default: return set_error(B200_ERR_INVALID_ARGUMENT, "allreduce: world size must be 2, 4 or 8, got %d", p.world);
  }
  return check_launch("allreduce_oneshot_kernel");
}


// ---- fused: all-reduce + residual add + RMSNorm ------------------------------------------------------------
// The two sums of a tensor-parallel decoder layer are each followed by `h += sum; x = rms_norm(h) * w`
// (reference layer wiring; SURVEY.md 8(e) asks for the fused form).  CTA b owns ROWS b, b + grid, ... on every
// rank (a row of the headline shape is 8 KiB: one 16-byte vector per thread), so after the same flag exchange as
// above it gathers a whole row from every peer, adds in rank order, rounds the sum to T exactly as the unfused
// all-reduce stores it, adds the residual (rounded to T, as b200_rms_norm does), keeps the row in registers for
// the sum of squares, and writes the residual stream and the normalised row -- one launch and one HBM round trip
// of [M, hidden] less than allreduce_oneshot_kernel + rms_norm_vec_kernel, bit-identical to that sequence.
struct FusedParams {
  Params ar;
  void* residual;       // [rows, hidden] in place
  const void* weight;   // [hidden] or null
  void* out;            // [rows, hidden] T, or fp16 when out_f16
  int rows, hidden;
  float eps;
  int out_f16;
  int twoshot;  // rows are reduced + normed by their OWNER rank (row % world) only and the results gathered
};

__device__ __forceinline__ float block_sum_512(float v, float* red) {
  v = warp_sum(v);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (lane == 0) red[warp] = v;
  __syncthreads();
  float t = lane < (kThreads >> 5) ? red[lane] : 0.f;
  t = warp_sum(t);
  __syncthreads();
  return t;
}

// TWOSHOT (8 ranks): the one-shot form makes every rank pull every row from 7 peers (3.5 MiB in per call at the
// headline shape) and norm all 64 rows 8 times over.  Here row r belongs to rank r % WORLD: the owner gathers it,
// adds, norms and publishes the new residual row and the normalised row in a second staging area; after a second
// flag exchange everybody copies the rows it does not own from their owners (0.45 + 0.9 MiB in).  Same arithmetic
// on the owner, plain copies elsewhere: bit-identical to the one-shot form on every rank.
// Staging per rank: [2 parities][half_bytes] partials | [2 parities][2: residual, out][half_bytes] results;
// flags: [grid][world] exchange 1 | [grid][world] exchange 2.
template <typename T, int WORLD, bool TWOSHOT>
__global__ void __launch_bounds__(kThreads) allreduce_add_rmsnorm_kernel(const FusedParams f) {
  pdl_trigger();
  pdl_wait();
  constexpr int NV = 2;  // vectors per thread: hidden <= 512 * 8 * 2
  __shared__ float red[32];
  const Params& p = f.ar;
  const int b = blockIdx.x, G = gridDim.x;
  const uint32_t e = p.epoch[b] + 1u;
  const int64_t half_off = (int64_t)(e & 1u) * p.half_bytes;
  const int nvec = f.hidden >> 3;  // 16-byte vectors per row
  uint4* xv = reinterpret_cast<uint4*>(p.x);
  uint4* mine = reinterpret_cast<uint4*>(p.symm[p.rank] + half_off);
  // 1. publish my rows
  for (int r = b; r < f.rows; r += G)
    for (int v = threadIdx.x; v < nvec; v += kThreads) mine[(int64_t)r * nvec + v] = xv[(int64_t)r * nvec + v];
  __syncthreads();
  // 2. + 3. flag exchange (same slots and epochs as allreduce_oneshot_kernel: the two kernels interleave freely)
  if (threadIdx.x < WORLD && (int)threadIdx.x != p.rank) {
    const int peer = threadIdx.x;
    st_release_sys(p.flags[peer] + (size_t)b * WORLD + p.rank, e);
    const uint32_t* slot = p.flags[p.rank] + (size_t)b * WORLD + peer;
    uint32_t spins = 0;
    while ((int32_t)(ld_acquire_sys(slot) - e) < 0) {
      if (++spins > (1u << 27)) __trap();
    }
  }
  __syncthreads();
  // 4. per row: gather + sum, residual add, norm
  uint4* rv = reinterpret_cast<uint4*>(f.residual);
  uint4* ov = reinterpret_cast<uint4*>(f.out);
  const uint4* wv = reinterpret_cast<const uint4*>(f.weight);
  const int64_t res_off = 2 * p.half_bytes + (int64_t)(e & 1u) * 2 * p.half_bytes;  // this call's result area
  uint4* my_h = reinterpret_cast<uint4*>(p.symm[p.rank] + res_off);
  uint4* my_o = reinterpret_cast<uint4*>(p.symm[p.rank] + res_off + p.half_bytes);
  for (int r = b; r < f.rows; r += G) {
    if (TWOSHOT && (r % WORLD) != p.rank) continue;
    float h[NV][8];
    float ss = 0.f;
#pragma unroll
    for (int j = 0; j < NV; ++j) {
      const int v = threadIdx.x + j * kThreads;
      if (v < nvec) {
        const int64_t idx = (int64_t)r * nvec + v;
        uint4 part[WORLD];
#pragma unroll
        for (int q = 0; q < WORLD; ++q)
          part[q] = (q == p.rank) ? xv[idx] : ld_peer_u4(reinterpret_cast<const uint4*>(p.symm[q] + half_off) + idx);
        float acc[8];
#pragma unroll
        for (int i = 0; i < 8; ++i) acc[i] = 0.f;
#pragma unroll
        for (int q = 0; q < WORLD; ++q) accumulate<T>(acc, part[q]);
        const uint4 res = rv[idx];
        const T* re = reinterpret_cast<const T*>(&res);
        uint4 hn;
        T* he = reinterpret_cast<T*>(&hn);
#pragma unroll
        for (int i = 0; i < 8; ++i) {
          const float a = to_float(from_float<T>(acc[i]));        // the all-reduce's own rounding
          he[i] = from_float<T>(a + to_float(re[i]));             // residual stream kept in T
          h[j][i] = to_float(he[i]);
          ss += h[j][i] * h[j][i];
        }
        rv[idx] = hn;
        if (TWOSHOT) my_h[idx] = hn;
      }
    }
    ss = block_sum_512(ss, red);
    const float inv = rsqrtf(ss / (float)f.hidden + f.eps);
#pragma unroll
    for (int j = 0; j < NV; ++j) {
      const int v = threadIdx.x + j * kThreads;
      if (v < nvec) {
        uint4 c;
        T* ce = reinterpret_cast<T*>(&c);
        if (wv) {
          const uint4 w = __ldg(wv + v);
          const T* we = reinterpret_cast<const T*>(&w);
#pragma unroll
          for (int i = 0; i < 8; ++i) ce[i] = from_float<T>(h[j][i] * inv * to_float(we[i]));
        } else {
#pragma unroll
          for (int i = 0; i < 8; ++i) ce[i] = from_float<T>(h[j][i] * inv);
        }
        if (f.out_f16) {
          __half* hh = reinterpret_cast<__half*>(&c);
#pragma unroll
          for (int i = 0; i < 8; ++i) hh[i] = __float2half_rn(fminf(fmaxf(to_float(ce[i]), -65504.f), 65504.f));
        }
        ov[(int64_t)r * nvec + v] = c;
        if (TWOSHOT) my_o[(int64_t)r * nvec + v] = c;
      }
    }
  }
  if constexpr (TWOSHOT) {
    __syncthreads();
    // 5. second exchange: my owned rows of this CTA's slice are published
    if (threadIdx.x < WORLD && (int)threadIdx.x != p.rank) {
      const int peer = threadIdx.x;
      st_release_sys(p.flags[peer] + (size_t)(G + b) * WORLD + p.rank, e);
      const uint32_t* slot = p.flags[p.rank] + (size_t)(G + b) * WORLD + peer;
      uint32_t spins = 0;
      while ((int32_t)(ld_acquire_sys(slot) - e) < 0) {
        if (++spins > (1u << 27)) __trap();
      }
    }
    __syncthreads();
    // 6. copy the rows other ranks own
    for (int r = b; r < f.rows; r += G) {
      const int owner = r % WORLD;
      if (owner == p.rank) continue;
      const uint4* oh = reinterpret_cast<const uint4*>(p.symm[owner] + res_off);
      const uint4* oo = reinterpret_cast<const uint4*>(p.symm[owner] + res_off + p.half_bytes);
      for (int v = threadIdx.x; v < nvec; v += kThreads) {
        const int64_t idx = (int64_t)r * nvec + v;
        const uint4 a = ld_peer_u4(oh + idx), c = ld_peer_u4(oo + idx);
        rv[idx] = a;
        ov[idx] = c;
      }
    }
  }
  __syncthreads();
  if (threadIdx.x == 0) p.epoch[b] = e;
}

template <typename T>
static int launch_fused(const FusedParams& f, int grid, cudaStream_t st) {
  switch (f.ar.world) {
    case 2:
      if (f.twoshot) B200_CUDA_OK(launch_pdl(allreduce_add_rmsnorm_kernel<T, 2, true>, dim3(grid), dim3(kThreads), 0, st, f));
      else B200_CUDA_OK(launch_pdl(allreduce_add_rmsnorm_kernel<T, 2, false>, dim3(grid), dim3(kThreads), 0, st, f));
      break;
    case 4:
      if (f.twoshot) B200_CUDA_OK(launch_pdl(allreduce_add_rmsnorm_kernel<T, 4, true>, dim3(grid), dim3(kThreads), 0, st, f));
      else B200_CUDA_OK(launch_pdl(allreduce_add_rmsnorm_kernel<T, 4, false>, dim3(grid), dim3(kThreads), 0, st, f));
      break;
    case 8:
      if (f.twoshot) B200_CUDA_OK(launch_pdl(allreduce_add_rmsnorm_kernel<T, 8, true>, dim3(grid), dim3(kThreads), 0, st, f));
      else B200_CUDA_OK(launch_pdl(allreduce_add_rmsnorm_kernel<T, 8, false>, dim3(grid), dim3(kThreads), 0, st, f));
      break;
    default: return set_error(B200_ERR_INVALID_ARGUMENT, "allreduce: world size must be 2, 4 or 8, got %d", f.ar.world);
  }
  return check_launch("allreduce_add_rmsnorm_kernel");
}

// ---- push form: reduce-scatter + norm on the row's owner, all-gather of the normed rows (LL packets) ------------
// The pull kernels above pay, per call, a release fence + a flag store + the peers' acquire polls + NVLink
// round-trip loads (24 us at 8 ranks, where the 3.5 MiB of pulls is also latency-, not bandwidth-, limited).  Here
// nothing is pulled and nothing is fenced.  Row r belongs to rank r % WORLD, which is also the only rank that keeps
// the residual stream of that row (the stream is row-sharded between the fused calls of a forward pass; every rank
// still gets every NORMED row, which is all the linears that follow consume):
//   phase 1  a rank that does not own row r stores its partial of r straight into the owner's inbox
//   owner    polls its inbox (local memory) until the W-1 partials of r are there, adds them to its own in RANK
//            ORDER in fp32, rounds to T, adds the residual, norms -- the arithmetic of the kernels above, bit for
//            bit -- and stores the normed row into every peer's outbox
//   phase 2  the other ranks poll their outbox (local memory) and copy the row to `out`
// Every store is an LL packet: 16 bytes = two 8-byte words {4 bytes of data, the call's epoch}.  A naturally
// aligned 8-byte scalar word is single-copy atomic, so a word whose upper half shows the epoch carries this call's
// data: the data is its own flag.  Wire bytes are 2x the payload, but the payload is 1/4 of the pull form's at 8
// ranks (a rank sends 7/8 of [M, hidden] once as partials and 1/8 of it seven times as results) and the latency
// chain is two one-way NVLink trips.  Inbox / outbox alternate by epoch parity; a rank can be at most one call ahead
// of a peer (it needs the peer's phase-2 rows of call n to leave call n, and the peer sends those only after it
// has consumed every phase-1 packet of call n), so a packet still being polled is never overwritten.
struct LLParams {
  const void* x;         // [rows, hidden] this rank's partial
  void* residual;        // [rows, hidden] in place; only rows this rank owns are read and written
  const void* weight;    // [hidden] or null
  void* out;             // [rows, hidden]
  char* box[kMaxWorld];  // every rank's packet buffer (peer-mapped): [2 parities][inbox | outbox], `phase_bytes` each
  uint32_t* epoch;       // local [grid]
  int64_t phase_bytes;
  int rows, hidden, rank, world;
  float eps;
  int out_f16;
};

__device__ __forceinline__ void ll_store(char* dst, const uint4& v, uint32_t e) {
  const uint64_t f = (uint64_t)e << 32;
  asm volatile("st.relaxed.sys.global.v2.u64 [%0], {%1,%2};" ::"l"(dst), "l"(f | v.x), "l"(f | v.y) : "memory");
  asm volatile("st.relaxed.sys.global.v2.u64 [%0], {%1,%2};" ::"l"(dst + 16), "l"(f | v.z), "l"(f | v.w) : "memory");
}
// one probe of a 32-byte packet pair; true when all four words carry epoch e
__device__ __forceinline__ bool ll_probe(const char* src, uint32_t e, uint4& v) {
  uint64_t a, b, c, d;
  asm volatile("ld.relaxed.sys.global.v2.u64 {%0,%1}, [%2];" : "=l"(a), "=l"(b) : "l"(src) : "memory");
  asm volatile("ld.relaxed.sys.global.v2.u64 {%0,%1}, [%2];" : "=l"(c), "=l"(d) : "l"(src + 16) : "memory");
  v = make_uint4((uint32_t)a, (uint32_t)b, (uint32_t)c, (uint32_t)d);
  return (uint32_t)(a >> 32) == e && (uint32_t)(b >> 32) == e && (uint32_t)(c >> 32) == e && (uint32_t)(d >> 32) == e;
}

// Work split: every CTA of every rank takes an equal share of all three steps, so no single SM has to issue a whole
// row's packets to W-1 peers (a first version with one CTA per owned row spent its time there: 15.4 us at 8 ranks).
// Clusters of W CTAs own a row together: cluster CTA c runs threads [c * 512 / W, (c + 1) * 512 / W) of the pull kernel's
// 512-thread row mapping (same vectors per thread, same warp butterflies), the 16 warp sums of squares meet in
// every CTA's shared memory through DSMEM and are added in the pull kernel's order -- the same bits.
// The cluster size is the world size: with rows == grid every rank owns grid / W rows = its number of clusters, one
// row per cluster at 2, 4 and 8 ranks alike (a fixed 8 made 2-rank runs walk 4 rows per cluster: 14.6 us vs 9.3).
constexpr int kPushThreads = 256;

template <typename T, int WORLD>
__global__ void __cluster_dims__(WORLD, 1, 1) __launch_bounds__(kPushThreads)
allreduce_add_rmsnorm_ll_kernel(const LLParams f) {
  constexpr int kPushCluster = WORLD;
  constexpr int kPushOwn = kThreads / kPushCluster;  // threads of a CTA that do the owner step (256 / 128 / 64)
  static_assert(kPushOwn <= kPushThreads && kPushOwn % 32 == 0, "owner threads");
  pdl_trigger();
  pdl_wait();
  constexpr int NV = 2;  // vectors per (pull-kernel) thread: hidden <= 512 * 8 * 2
  __shared__ float red[2][kThreads / 32];
  const int b = blockIdx.x, G = gridDim.x;
  const int c = b % kPushCluster, cl = b / kPushCluster, ncl = G / kPushCluster;  // clusters span x: c = cluster rank
  const uint32_t e = f.epoch[b] + 1u;
  const int nvec = f.hidden >> 3;
  const int64_t par_off = (int64_t)(e & 1u) * 2 * f.phase_bytes;
  const uint4* xv = reinterpret_cast<const uint4*>(f.x);
  uint4* ov = reinterpret_cast<uint4*>(f.out);
  const int64_t total = (int64_t)f.rows * nvec;
  // inbox slot of (row r, source s) on r's owner: packet row (r / WORLD) * WORLD + s; outbox slot of row r: packet row r

  // ---- phase 1: my partials of the rows I do not own go to their owners ----
  for (int64_t j = (int64_t)b * kPushThreads + threadIdx.x; j < total; j += (int64_t)G * kPushThreads) {
    const int r = (int)(j / nvec), v = (int)(j - (int64_t)r * nvec);
    const int o = r % WORLD;
    if (o == f.rank) continue;
    const int64_t slot = ((int64_t)(r / WORLD) * WORLD + f.rank) * nvec + v;
    ll_store(f.box[o] + par_off + slot * 32, xv[j], e);
  }

  // ---- owner step: rows rank, rank + W, ...; one cluster per row ----
  const char* inbox = f.box[f.rank] + par_off;
  uint4* rv = reinterpret_cast<uint4*>(f.residual);
  const uint4* wv = reinterpret_cast<const uint4*>(f.weight);
  const int n_owned = f.rows > f.rank ? (f.rows - f.rank + WORLD - 1) / WORLD : 0;
  const bool own_thread = threadIdx.x < kPushOwn;
  const int ot = c * kPushOwn + threadIdx.x;  // thread id in the pull kernel's row mapping
  int it = 0;
  for (int i = cl; i < n_owned; i += ncl, ++it) {
    const int r = f.rank + i * WORLD;
    float h[NV][8];
    float ss = 0.f;
    if (own_thread) {
#pragma unroll
      for (int j = 0; j < NV; ++j) {
        const int v = ot + j * kThreads;
        if (v < nvec) {
          const int64_t idx = (int64_t)r * nvec + v;
          uint4 part[WORLD];
          const uint4 own = xv[idx];
          const int64_t slot0 = (int64_t)(r / WORLD) * WORLD * nvec + v;
          uint32_t spins = 0;
          bool all;
          do {  // all W-1 packets are probed in one batch: one local-memory latency per round, not W-1
            all = true;
#pragma unroll
            for (int q = 0; q < WORLD; ++q) {
              uint4 t = own;
              const bool okq = q == f.rank || ll_probe(inbox + (slot0 + (int64_t)q * nvec) * 32, e, t);
              part[q] = t;
              all = all && okq;
            }
            if (++spins > (1u << 27)) __trap();  // a dead peer must fault this rank, not hang the box
          } while (!all);
          float acc[8];
#pragma unroll
          for (int k = 0; k < 8; ++k) acc[k] = 0.f;
#pragma unroll
          for (int q = 0; q < WORLD; ++q) accumulate<T>(acc, part[q]);
          const uint4 res = rv[idx];
          const T* re = reinterpret_cast<const T*>(&res);
          uint4 hn;
          T* he = reinterpret_cast<T*>(&hn);
#pragma unroll
          for (int k = 0; k < 8; ++k) {
            const float a = to_float(from_float<T>(acc[k]));  // the all-reduce's own rounding
            he[k] = from_float<T>(a + to_float(re[k]));       // residual stream kept in T
            h[j][k] = to_float(he[k]);
            ss += h[j][k] * h[j][k];
          }
          rv[idx] = hn;
        }
      }
      // block_sum_512 of the pull kernel, spread over the cluster: warp butterfly, 16 warp sums into every CTA
      ss = warp_sum(ss);
      if ((threadIdx.x & 31) == 0) {
        const uint32_t local = (uint32_t)__cvta_generic_to_shared(&red[it & 1][ot >> 5]);
#pragma unroll
        for (int q = 0; q < kPushCluster; ++q) {
          uint32_t remote;
          asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(remote) : "r"(local), "r"(q));
          asm volatile("st.shared::cluster.f32 [%0], %1;" ::"r"(remote), "f"(ss) : "memory");
        }
      }
    }
    asm volatile("barrier.cluster.arrive.release.aligned;" ::: "memory");
    asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory");
    if (own_thread) {
      const int lane = threadIdx.x & 31;
      float t = lane < (kThreads >> 5) ? red[it & 1][lane] : 0.f;
      t = warp_sum(t);
      const float inv = rsqrtf(t / (float)f.hidden + f.eps);
#pragma unroll
      for (int j = 0; j < NV; ++j) {
        const int v = ot + j * kThreads;
        if (v < nvec) {
          uint4 cvec;
          T* ce = reinterpret_cast<T*>(&cvec);
          if (wv) {
            const uint4 w = __ldg(wv + v);
            const T* we = reinterpret_cast<const T*>(&w);
#pragma unroll
            for (int k = 0; k < 8; ++k) ce[k] = from_float<T>(h[j][k] * inv * to_float(we[k]));
          } else {
#pragma unroll
            for (int k = 0; k < 8; ++k) ce[k] = from_float<T>(h[j][k] * inv);
          }
          if (f.out_f16) {
            __half* hh = reinterpret_cast<__half*>(&cvec);
#pragma unroll
            for (int k = 0; k < 8; ++k) hh[k] = __float2half_rn(fminf(fmaxf(to_float(ce[k]), -65504.f), 65504.f));
          }
          const int64_t idx = (int64_t)r * nvec + v;
#pragma unroll
          for (int q = 0; q < WORLD; ++q)
            if (q != f.rank) ll_store(f.box[q] + par_off + f.phase_bytes + idx * 32, cvec, e);
          ov[idx] = cvec;
        }
      }
    }
  }

  // ---- phase 2: the normed rows of the other owners arrive in my outbox ----
  const char* outbox = f.box[f.rank] + par_off + f.phase_bytes;
  for (int64_t j = (int64_t)b * kPushThreads + threadIdx.x; j < total; j += (int64_t)G * kPushThreads) {
    const int r = (int)(j / nvec);
    if (r % WORLD == f.rank) continue;
    uint4 cvec;
    uint32_t spins = 0;
    while (!ll_probe(outbox + j * 32, e, cvec)) {
      if (++spins > (1u << 27)) __trap();
    }
    ov[j] = cvec;
  }
  __syncthreads();
  if (threadIdx.x == 0) f.epoch[b] = e;
}

template <typename T>
static int launch_ll(const LLParams& f, int grid, cudaStream_t st) {
  switch (f.world) {
    case 2: B200_CUDA_OK(launch_pdl(allreduce_add_rmsnorm_ll_kernel<T, 2>, dim3(grid), dim3(kPushThreads), 0, st, f)); break;
    case 4: B200_CUDA_OK(launch_pdl(allreduce_add_rmsnorm_ll_kernel<T, 4>, dim3(grid), dim3(kPushThreads), 0, st, f)); break;
    case 8: B200_CUDA_OK(launch_pdl(allreduce_add_rmsnorm_ll_kernel<T, 8>, dim3(grid), dim3(kPushThreads), 0, st, f)); break;
    default: return set_error(B200_ERR_INVALID_ARGUMENT, "allreduce: world size must be 2, 4 or 8, got %d", f.world);
  }
  return check_launch("allreduce_add_rmsnorm_ll_kernel");
}

// ---- vocab-parallel greedy head: (max logit, token) exchange (lm_head row) ----------------
// Every rank holds the logits of its own vocab slice; the greedy token is the arg-max over all slices.  Each rank
// PUSHES its per-row winner (value, global index) into every peer's slot table with two 8-byte stores that carry
// the call's epoch in their upper word (NCCL's LL idea: the data is its own flag, no separate signal and no fence),
// then polls its own table -- local memory -- until all `world` entries of the row show the epoch.  One NVLink
// one-way latency per call.  Slots alternate by epoch parity: a rank can run at most one call ahead of a peer (it
// needs the peer's entry of call n to leave call n), so the entry a slow rank is still polling is never overwritten.
// Ties go to the lowest global index, exactly as one arg-max over the gathered logits would (mx.argmax).
struct ArgmaxXParams {
  const void* logits;
  int32_t* out;       // in: local arg-max per row (b200_argmax over the slice); out: global token id
  uint2* slots[kMaxWorld];
  uint32_t* epoch;    // local uint32 [max_rows]
  int64_t stride;
  int rows, max_rows, vocab_offset, rank, world;
};

__device__ __forceinline__ void st_relaxed_sys_v2(uint2* p, uint32_t a, uint32_t b) {
  asm volatile("st.relaxed.sys.global.v2.u32 [%0], {%1,%2};" ::"l"(p), "r"(a), "r"(b) : "memory");
}
__device__ __forceinline__ uint2 ld_relaxed_sys_v2(const uint2* p) {
  uint2 v;
  asm volatile("ld.relaxed.sys.global.v2.u32 {%0,%1}, [%2];" : "=r"(v.x), "=r"(v.y) : "l"(p) : "memory");
  return v;
}

template <typename T>
__global__ void __launch_bounds__(32) argmax_exchange_kernel(const ArgmaxXParams p) {
  pdl_trigger();
  pdl_wait();
  const int r = blockIdx.x, lane = threadIdx.x;
  const uint32_t e = p.epoch[r] + 1;
  const int li = p.out[r];
  const float v = to_float(reinterpret_cast<const T*>(p.logits)[(int64_t)r * p.stride + li]);
  const int gi = p.vocab_offset + li;
  __syncwarp();
  float best = -INFINITY;
  int bi = 0x7fffffff;
  if (lane < p.world) {
    const size_t half = (size_t)(e & 1u) * p.world;
    uint2* dst = p.slots[lane] + ((half + p.rank) * p.max_rows + r) * 2;  // entry [parity][this rank][row] on rank `lane`
    st_relaxed_sys_v2(dst, __float_as_uint(v), e);
    st_relaxed_sys_v2(dst + 1, (uint32_t)gi, e);
    const uint2* src = p.slots[p.rank] + ((half + lane) * p.max_rows + r) * 2;  // rank `lane`'s entry in my table
    uint2 a, b;
    uint32_t spins = 0;
    do {
      a = ld_relaxed_sys_v2(src);
      b = ld_relaxed_sys_v2(src + 1);
      if (++spins > (1u << 27)) __trap();  // a dead peer must fault this rank, not hang the box
    } while (a.y != e || b.y != e);
    best = __uint_as_float(a.x);
    bi = (int)b.x;
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    const float ov = __shfl_xor_sync(0xffffffffu, best, o);
    const int oi = __shfl_xor_sync(0xffffffffu, bi, o);
    if (ov > best || (ov == best && oi < bi)) {
      best = ov;
      bi = oi;
    }
  }
  if (lane == 0) {
    p.out[r] = bi == 0x7fffffff ? 0 : bi;
    p.epoch[r] = e;
  }
}

}  // namespace ar
}  // namespace b200

extern "C" {

int b200_allreduce_add_rmsnorm_push(const void* x, void* residual, const void* weight, void* out, int rows, int hidden,
                                    float eps, int dtype, int out_dtype, void* const* box_ptrs, int64_t box_bytes,
                                    void* epoch, int rank, int world, int grid, void* stream) {
  using namespace b200;
  if (rows == 0) return B200_OK;
  B200_CHECK_ARG(x && residual && out && box_ptrs && epoch && rows > 0, "allreduce_add_rmsnorm_push: bad arguments");
  B200_CHECK_ARG(dtype == B200_BF16 || dtype == B200_F16, "allreduce_add_rmsnorm_push: dtype must be bf16 or fp16");
  B200_CHECK_ARG(out_dtype == dtype || (out_dtype == B200_F16 && dtype == B200_BF16),
                 "allreduce_add_rmsnorm_push: out dtype must be the input dtype, or fp16 for bf16 input");
  B200_CHECK_ARG(hidden % 8 == 0 && hidden <= ar::kThreads * 8 * 2,
                 "allreduce_add_rmsnorm_push: hidden %d must be a multiple of 8, <= %d", hidden, ar::kThreads * 8 * 2);
  B200_CHECK_ARG(rank >= 0 && rank < world && world <= ar::kMaxWorld, "allreduce: rank %d of world %d", rank, world);
  B200_CHECK_ARG(grid > 0 && grid <= 1024 && grid % world == 0,
                 "allreduce_add_rmsnorm_push: grid %d must be a positive multiple of the world size %d (CTAs per cluster)",
                 grid, world);
  B200_CHECK_ARG((((uintptr_t)x | (uintptr_t)residual | (uintptr_t)out | (uintptr_t)weight) & 15) == 0,
                 "allreduce_add_rmsnorm_push: pointers must be 16-byte aligned");
  const int64_t phase_bytes = (box_bytes / 4) & ~(int64_t)255;  // [2 parities][inbox | outbox]
  const int64_t packet_rows = ((int64_t)rows + world - 1) / world * world;
  B200_CHECK_ARG(packet_rows * hidden * 4 <= phase_bytes,
                 "allreduce_add_rmsnorm_push: %d rows x %d need %lld packet bytes per phase, the buffer has %lld", rows,
                 hidden, (long long)(packet_rows * hidden * 4), (long long)phase_bytes);
  ar::LLParams f{};
  f.x = x;
  f.residual = residual;
  f.weight = weight;
  f.out = out;
  for (int r = 0; r < world; ++r) f.box[r] = reinterpret_cast<char*>(box_ptrs[r]);
  f.epoch = reinterpret_cast<uint32_t*>(epoch);
  f.phase_bytes = phase_bytes;
  f.rows = rows;
  f.hidden = hidden;
  f.rank = rank;
  f.world = world;
  f.eps = eps;
  f.out_f16 = (out_dtype == B200_F16 && dtype == B200_BF16) ? 1 : 0;
  cudaStream_t st = (cudaStream_t)stream;
  if (dtype == B200_BF16) return ar::launch_ll<__nv_bfloat16>(f, grid, st);
  return ar::launch_ll<__half>(f, grid, st);
}

int b200_argmax_vocab_parallel(const void* logits, int32_t* out, int rows, int shard_vocab, int64_t stride, int dtype,
                               int vocab_offset, void* const* slot_ptrs, void* epoch, int rank, int world,
                               int max_rows, void* stream) {
  using namespace b200;
  if (rows == 0) return B200_OK;
  B200_CHECK_ARG(logits && out && slot_ptrs && epoch && rows > 0 && shard_vocab > 0, "argmax_vocab_parallel: bad arguments");
  B200_CHECK_ARG(rank >= 0 && rank < world && world >= 1 && world <= ar::kMaxWorld,
                 "argmax_vocab_parallel: rank %d of world %d", rank, world);
  B200_CHECK_ARG(rows <= max_rows, "argmax_vocab_parallel: %d rows exceed the slot table (%d)", rows, max_rows);
  int rc = b200_argmax(logits, out, rows, shard_vocab, stride, dtype, stream);  // winner of this rank's slice
  if (rc != B200_OK) return rc;
  ar::ArgmaxXParams p{};
  p.logits = logits;
  p.out = out;
  for (int r = 0; r < world; ++r) p.slots[r] = reinterpret_cast<uint2*>(slot_ptrs[r]);
  p.epoch = reinterpret_cast<uint32_t*>(epoch);
  p.stride = stride;
  p.rows = rows;
  p.max_rows = max_rows;
  p.vocab_offset = vocab_offset;
  p.rank = rank;
  p.world = world;
  cudaStream_t st = (cudaStream_t)stream;
  return B200_DISPATCH_DTYPE(dtype, T, [&]() -> int {
    launch_pdl(ar::argmax_exchange_kernel<T>, dim3(rows), dim3(32), 0, st, p);
    return check_launch("argmax_exchange_kernel");
  });
}

int b200_ipc_alloc(int64_t bytes, void** ptr, unsigned char* handle64) {
  using namespace b200;
  B200_CHECK_ARG(bytes > 0 && ptr && handle64, "ipc_alloc: bad arguments");
  static_assert(sizeof(cudaIpcMemHandle_t) == 64, "IPC handle is 64 bytes");
  void* d = nullptr;
  B200_CUDA_OK(cudaMalloc(&d, (size_t)bytes));
  B200_CUDA_OK(cudaMemset(d, 0, (size_t)bytes));
  B200_CUDA_OK(cudaDeviceSynchronize());
  cudaIpcMemHandle_t h;
  B200_CUDA_OK(cudaIpcGetMemHandle(&h, d));
  memcpy(handle64, &h, 64);
  *ptr = d;
  return B200_OK;
}

int b200_ipc_open(const unsigned char* handle64, void** ptr) {
  using namespace b200;
  B200_CHECK_ARG(ptr && handle64, "ipc_open: bad arguments");
  cudaIpcMemHandle_t h;
  memcpy(&h, handle64, 64);
  B200_CUDA_OK(cudaIpcOpenMemHandle(ptr, h, cudaIpcMemLazyEnablePeerAccess));
  return B200_OK;
}

int b200_ipc_close(void* ptr) {
  using namespace b200;
  B200_CUDA_OK(cudaIpcCloseMemHandle(ptr));
  return B200_OK;
}

int b200_ipc_free(void* ptr) {
  using namespace b200;
  B200_CUDA_OK(cudaFree(ptr));
  return B200_OK;
}

int b200_allreduce_sum(void* x, int64_t numel, int dtype, void* const* symm_ptrs, void* const* flag_ptrs,
                       void* epoch, int rank, int world, int64_t half_bytes, int grid, void* stream) {
  using namespace b200;
  if (numel == 0) return B200_OK;
  B200_CHECK_ARG(x && symm_ptrs && flag_ptrs && epoch && numel > 0, "allreduce: bad arguments");
  B200_CHECK_ARG(rank >= 0 && rank < world && world <= ar::kMaxWorld, "allreduce: rank %d of world %d", rank, world);
  B200_CHECK_ARG(grid > 0 && grid <= 1024, "allreduce: grid %d out of range", grid);
  const int64_t bytes = numel * (int64_t)dtype_size(dtype);
  B200_CHECK_ARG(bytes % 16 == 0 && ((uintptr_t)x & 15) == 0, "allreduce: x must be 16-byte aligned and a multiple of 16 bytes");
  B200_CHECK_ARG(bytes <= half_bytes, "allreduce: %lld bytes exceed the staging half (%lld)", (long long)bytes,
                 (long long)half_bytes);
  ar::Params p;
  p.x = x;
  for (int r = 0; r < world; ++r) {
    p.symm[r] = reinterpret_cast<char*>(symm_ptrs[r]);
    p.flags[r] = reinterpret_cast<uint32_t*>(flag_ptrs[r]);
  }
  p.epoch = reinterpret_cast<uint32_t*>(epoch);
  p.nvec = bytes / 16;
  p.half_bytes = half_bytes;
  p.rank = rank;
  p.world = world;
  cudaStream_t st = (cudaStream_t)stream;
  return B200_DISPATCH_DTYPE(dtype, T, [&]() -> int { return ar::launch<T>(p, grid, st); });
}

int b200_allreduce_add_rmsnorm(void* x, void* residual, const void* weight, void* out, int rows, int hidden, float eps,
                               int dtype, int out_dtype, void* const* symm_ptrs, void* const* flag_ptrs, void* epoch,
                               int rank, int world, int64_t half_bytes, int grid, int twoshot, void* stream) {
  using namespace b200;
  if (rows == 0) return B200_OK;
  B200_CHECK_ARG(x && residual && out && symm_ptrs && flag_ptrs && epoch && rows > 0, "allreduce_add_rmsnorm: bad arguments");
  B200_CHECK_ARG(dtype == B200_BF16 || dtype == B200_F16, "allreduce_add_rmsnorm: dtype must be bf16 or fp16");
  B200_CHECK_ARG(out_dtype == dtype || (out_dtype == B200_F16 && dtype == B200_BF16),
                 "allreduce_add_rmsnorm: out dtype must be the input dtype, or fp16 for bf16 input");
  B200_CHECK_ARG(hidden % 8 == 0 && hidden <= ar::kThreads * 8 * 2, "allreduce_add_rmsnorm: hidden %d must be a multiple of 8, <= %d",
                 hidden, ar::kThreads * 8 * 2);
  B200_CHECK_ARG(rank >= 0 && rank < world && world <= ar::kMaxWorld, "allreduce: rank %d of world %d", rank, world);
  B200_CHECK_ARG(grid > 0 && grid <= 1024, "allreduce: grid %d out of range", grid);
  const int64_t bytes = (int64_t)rows * hidden * 2;
  B200_CHECK_ARG((((uintptr_t)x | (uintptr_t)residual | (uintptr_t)out | (uintptr_t)weight) & 15) == 0,
                 "allreduce_add_rmsnorm: pointers must be 16-byte aligned");
  B200_CHECK_ARG(bytes <= half_bytes, "allreduce: %lld bytes exceed the staging half (%lld)", (long long)bytes,
                 (long long)half_bytes);
  ar::FusedParams f;
  f.ar.x = x;
  for (int r = 0; r < world; ++r) {
    f.ar.symm[r] = reinterpret_cast<char*>(symm_ptrs[r]);
    f.ar.flags[r] = reinterpret_cast<uint32_t*>(flag_ptrs[r]);
  }
  f.ar.epoch = reinterpret_cast<uint32_t*>(epoch);
  f.ar.nvec = bytes / 16;
  f.ar.half_bytes = half_bytes;
  f.ar.rank = rank;
  f.ar.world = world;
  f.residual = residual;
  f.weight = weight;
  f.out = out;
  f.rows = rows;
  f.hidden = hidden;
  f.eps = eps;
  f.out_f16 = (out_dtype == B200_F16 && dtype == B200_BF16) ? 1 : 0;
  f.twoshot = twoshot ? 1 : 0;
  cudaStream_t st = (cudaStream_t)stream;
  if (dtype == B200_BF16) return ar::launch_fused<__nv_bfloat16>(f, grid, st);
  return ar::launch_fused<__half>(f, grid, st);
}

}  // extern "C"
