// (a4) Paged GQA decode, warp-specialised persistent kernel fed by TMA (sm_100a).
//
// Same arithmetic as decode_kernel (decode.cu: G x D query block padded to the M = 16 of mma.sync.m16n8k16, each
// of four consumer warps takes 16 keys of a 64-key tile, fp32 online softmax) and the same flat tile-stream work
// partition as decode_sk_kernel (decode_sk.cu: the batch is ONE unit-major stream of 64-key tiles, CTA c of N owns
// the contiguous range [c*T/N, (c+1)*T/N), units cut across CTAs are merged by the last piece to arrive).  What
// changes is how bytes reach shared memory and who waits for them:
//
//   warp 4      PRODUCER.  Walks the CTA's tile range and keeps a 3-stage ring (32 KB per stage: K + V of 64
//               keys) full through the copy engine: per tile the warp reads the 64 slot ids and, when every aligned
//               16-key group sits in 16 consecutive pool slots (true under the paged allocator for page_size >= 16),
//               issues 16-row tiled TMA boxes (one instruction per 4 KB); otherwise - page_size 1, fragmented or
//               ragged last tiles - tile::gather4 (4 arbitrary pool rows per instruction, rows past the end
//               zero-filled).  It also brings each segment's G x D query block in with one bulk copy, and with
//               `new_k / new_v` it writes the step's new K/V row into the pool itself right before the tile that
//               reads it (the separate kv_store launch disappears).  The producer runs AHEAD ACROSS segment
//               boundaries: while the consumers merge and write a finished unit, the next unit's first tiles and
//               query block are already landing, so the three dependent DRAM latencies (kv_indptr -> slot ids ->
//               first K row) that every CTA of decode_kernel pays per unit are paid once per CTA.
//   warps 0-3   CONSUMERS.  mbarrier-wait on a stage, S = Q K^T and O += P V straight out of the 128B-swizzled
//               tiles (ldmatrix / .trans), release the stage.  At the end of a segment the four warps merge through
//               the LAST tile's stage (its 32 KB are exactly the [4][16][D] fp32 scratch the merge needs) before
//               releasing it, so no shared memory is reserved for the epilogue.
//
// Algorithmic bytes per (request, kv head): 2 * L * D * 2 (K and V rows) + 4 * L (slot ids).
#include "decode.cuh"
#include "sm100.cuh"

namespace b200attn {

constexpr int kWsThreads = 192;  // 4 consumer warps + 1 producer warp + 1 merger warp
constexpr int kWsMq = 4;         // consumer -> merger hand-off ring
constexpr int kWsConsumers = 128;
constexpr int kWsStages = 3;
constexpr int kWsMinChunk = 2;    // smallest chunk of the guided schedule (tiles)
constexpr int kWsMaxLevels = 12;  // schedule levels; chunks <= (kWsMaxLevels + 1) * CTAs

struct WsMsg {  // chunk index (-1: no more work), the request its first tile lies in, tiles of the requests before it
  int k, item;
  int64_t pre;
};

struct WsItem {
  int kv_beg, L, tiles;
};

__device__ __forceinline__ WsItem ws_item(const DecodeParams& p, int item) {
  const int req = p.req_ids ? p.req_ids[item] : item;
  int kv_beg = p.kv_indptr[req];
  int L = p.kv_indptr[req + 1] - kv_beg;
  const int lo = window_lo(L, p.window_left);
  WsItem it;
  it.kv_beg = kv_beg + lo;
  it.L = L - lo;
  it.tiles = max(1, (it.L + kDecTile - 1) / kDecTile);  // an empty request still owns one (fully masked) tile
  return it;
}

// CTA that owns global tile g under the partition start(c) = floor(c * T / N)
__device__ __forceinline__ int ws_owner(int64_t g, int64_t T, int N) {
  return (int)(((g + 1) * N - 1) / T);
}

template <int D>
struct WsSmem {
  static constexpr int kHalves = D / 64;                    // 64-column (128-byte) halves of a row
  static constexpr int kHalfBytes = kDecTile * 128;         // 64 rows x 128 B
  static constexpr int kTileBytes = kHalves * kHalfBytes;   // one operand, one stage
  static constexpr int kStageBytes = 2 * kTileBytes;        // K + V
  static constexpr int kRing = 0;
  static constexpr int kQBytes = kDecMaxG * D * 2;          // one segment's query block
  static constexpr int kQ = kRing + kWsStages * kStageBytes;
  static constexpr int kML = kQ + 2 * kQBytes;              // float m[4][16], l[4][16]
  static constexpr int kBar = kML + 512;
  static constexpr int kTotal = kBar + 8 * 24;
};

enum WsBar {
  kWsFull = 0,
  kWsEmpty = kWsStages,
  kWsQFull = 2 * kWsStages,
  kWsQEmpty = 2 * kWsStages + 2,
  kWsMqFull = 2 * kWsStages + 4,
  kWsMqEmpty = kWsMqFull + kWsMq,
  kWsBarCount = kWsMqEmpty + kWsMq
};
struct WsMerge {  // one partial piece of a unit, handed from the consumers to the merger warp
  int unit, qrow, h0, gcount, c_first, c_last;
};

// byte offset of 16-byte chunk `chunk` of row `row` inside one operand tile laid out by TMA with SWIZZLE_128B:
// [half][row][128 B], chunk index XOR (row & 7) inside each 128-byte line (tile base 1024-byte aligned)
template <int D>
__device__ __forceinline__ uint32_t ws_off(int row, int chunk) {
  return (uint32_t)((chunk >> 3) * (kDecTile * 128) + row * 128 + (((chunk & 7) ^ (row & 7)) << 4));
}

__device__ __forceinline__ void bulk_load_1d(uint32_t dst, const void* src, uint32_t bytes, uint32_t bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];\n" ::"r"(dst),
               "l"(src), "r"(bytes), "r"(bar)
               : "memory");
}
__device__ __forceinline__ void consumer_sync() { asm volatile("bar.sync 1, 128;\n" ::: "memory"); }

// dev-only timeline (make trace: -DWS_TRACE): p.lse is taken over as int64 [CTAs][8] of %globaltimer stamps
#ifdef WS_TRACE
__device__ __forceinline__ void ws_stamp(const DecodeParams& p, int slot) {
  long long t;
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
  reinterpret_cast<long long*>(p.lse)[(int64_t)blockIdx.x * 8 + slot] = t;
}
#define WS_STAMP(cond, slot) if (cond) ws_stamp(p, slot)
#define WS_LSE_ON false
#else
#define WS_STAMP(cond, slot)
#define WS_LSE_ON (p.lse != nullptr)
#endif

struct WsFused {  // the step's new K/V rows, written into the pool by the producer (nullptr: the caller stored them)
  const void* new_k;
  const void* new_v;
  int64_t new_stride;  // elements between token rows of new_k / new_v
  int skip_slot;       // rows whose slot equals this are not written (the reserved padding slot); -1: write all
};

template <int D, bool kBf16>
__global__ void __launch_bounds__(kWsThreads, 2)
    decode_ws_kernel(const __grid_constant__ CUtensorMap tm_kg, const __grid_constant__ CUtensorMap tm_vg,
                     const __grid_constant__ CUtensorMap tm_kb, const __grid_constant__ CUtensorMap tm_vb,
                     const DecodeParams p, const WsFused fused, const int min_tiles, const int num_slots,
                     const int ws_rows) {
  using namespace sm100;
  using T = typename std::conditional<kBf16, __nv_bfloat16, __half>::type;
  using S = WsSmem<D>;
  constexpr int kHalves = S::kHalves;
  extern __shared__ uint8_t smem_raw[];
  __shared__ int64_t s_red[kWsThreads / 32];
  __shared__ int64_t s_loff[kWsMaxLevels + 1];  // first tile of every schedule level (+ total)
  __shared__ int s_lfirst[kWsMaxLevels + 1];    // first chunk index of every level (+ number of chunks)
  __shared__ int s_lsize[kWsMaxLevels];         // chunk size (tiles) of every level
  __shared__ int s_nlev;
  __shared__ WsMsg s_msg[2];                    // producer -> consumers: the chunk a segment opens
  __shared__ WsMerge s_mq[kWsMq];

  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int HG = p.Hkv * p.groups_per_kvh;
  WS_STAMP(tid == 0, 0);
  // programmatic dependent launch: everything above ran while the previous kernel in the stream drained; global
  // memory written by it (the pool, kv_indptr / kv_indices) is only read after this wait.  No-op when the kernel
  // was launched without the attribute.
  asm volatile("griddepcontrol.wait;\n" ::: "memory");
  WS_STAMP(tid == 0, 1);

  // ---- work list: the batch is ONE unit-major stream of 64-key tiles,
  //     g = HG * (tiles of the requests before i) + hg * tiles_i + t        (unit = (request i, kv head group hg)),
  // cut into CHUNKS by a guided schedule that every CTA derives for itself: level 0 hands each of the N CTAs one
  // chunk of ~ half its fair share (static: chunk = blockIdx.x, no atomic), every further level halves the chunk
  // size over what is left, down to kWsMinChunk tiles; chunks beyond level 0 are drawn from a global ticket by
  // whichever CTA runs dry first.  SMs do not stream at the same rate (measured: equal static ranges finish
  // between 144 and 177 us on cfg2), so equal ranges leave a ~7 % tail; with tickets the tail is one small chunk.
  // Chunk boundaries are a pure function of (total, N): pieces of a unit are found arithmetically for the merge.
  int64_t mine = 0;
  for (int i = tid; i < p.B; i += kWsThreads) mine += ws_item(p, i).tiles;
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) mine += __shfl_xor_sync(0xffffffffu, mine, o);
  if (lane == 0) s_red[warp] = mine;
  __syncthreads();
  int64_t total = 0;
#pragma unroll
  for (int w = 0; w < kWsThreads / 32; ++w) total += s_red[w];
  total *= HG;
  const int N = (int)gridDim.x;
  if (tid == 0) {
    const int64_t cmin = max(kWsMinChunk, min_tiles);
    int64_t off = 0, first = 0;
    int nl = 0;
    while (off < total) {
      const int64_t R = total - off;
      const bool last = (nl == kWsMaxLevels - 1);
      const int64_t S = max(cmin, (R + 2 * N - 1) / (2 * N));
      const int64_t all = (R + S - 1) / S;
      const int64_t cnt = (S == cmin || last) ? all : min((int64_t)N, all);
      s_loff[nl] = off;
      s_lfirst[nl] = (int)first;
      s_lsize[nl] = (int)S;
      off = min(total, off + cnt * S);
      first += cnt;
      ++nl;
    }
    s_loff[nl] = total;
    s_lfirst[nl] = (int)first;
    s_nlev = nl;
  }
  const uint32_t smem0 = (smem_u32(smem_raw) + 1023u) & ~1023u;
  uint8_t* smem = smem_raw + (smem0 - smem_u32(smem_raw));
  const uint32_t bar0 = smem0 + S::kBar;
  auto bar = [&](int i) { return bar0 + 8u * i; };
  if (tid == 0) {
    for (int s = 0; s < kWsStages; ++s) {
      mbar_init(bar(kWsFull + s), 1);
      mbar_init(bar(kWsEmpty + s), 4);  // one release per consumer warp
    }
    for (int s = 0; s < 2; ++s) {
      mbar_init(bar(kWsQFull + s), 1);
      mbar_init(bar(kWsQEmpty + s), 4);
    }
    for (int s = 0; s < kWsMq; ++s) {
      mbar_init(bar(kWsMqFull + s), 1);
      mbar_init(bar(kWsMqEmpty + s), 1);
    }
    mbar_fence_init();
  }
  __syncthreads();
  asm volatile("griddepcontrol.launch_dependents;\n" ::: "memory");
  WS_STAMP(tid == 0, 2);

  // ---- schedule arithmetic (shared memory tables, read-only from here on) ----------------------------------------
  const int nlev = s_nlev;
  const int num_chunks = s_lfirst[nlev];
  auto chunk_bounds = [&](int k, int64_t& b, int64_t& e) {
    int L = 0;
    while (k >= s_lfirst[L + 1]) ++L;
    b = s_loff[L] + (int64_t)(k - s_lfirst[L]) * s_lsize[L];
    e = min(b + s_lsize[L], s_loff[L + 1]);
  };
  auto chunk_of = [&](int64_t g) -> int {  // chunk that holds global tile g
    int L = 0;
    while (g >= s_loff[L + 1]) ++L;
    return s_lfirst[L] + (int)((g - s_loff[L]) / s_lsize[L]);
  };
  int* sched_next = p.ws_cnt + kDecMaxUnits - 2;  // ticket of the dynamically drawn chunks
  int* sched_done = p.ws_cnt + kDecMaxUnits - 1;  // CTAs that have drawn their last ticket
  const int num_static = min((int)gridDim.x, s_lfirst[nlev > 0 ? 1 : 0]);  // level 0: chunk = blockIdx.x

  // ---- segment walk state (both roles) --------------------------------------------------------------------------
  int item = 0;
  int64_t pre = 0;  // tiles of the requests before `item`
  WsItem it;
  int64_t cur = 0, g1 = 0;  // position in / end of the current chunk
  int n = 0;     // tiles produced / consumed so far (ring position)
  int nseg = 0;  // segments so far (query-block buffer parity)

  if (warp == 4) {
    // ================================ producer ================================
    if (lane == 0) {
      tma_prefetch_desc(&tm_kg);
      tma_prefetch_desc(&tm_vg);
      tma_prefetch_desc(&tm_kb);
      tma_prefetch_desc(&tm_vb);
    }
    // Flat, software-pipelined walk over the CTA's tiles: the slot ids (and the new-token slot) of tile n + 1 are
    // requested BEFORE tile n is issued, across segment and request boundaries, so the producer never sits on a
    // dependent index load between two TMA issues.
    struct Tile {
      const int32_t* idx;
      int key0, k_hi, col, kvh, item, L;
      bool last_of_unit;
    };
    int t = 0, t_hi = 0, hg = 0;
    int64_t ue = 0, seg_end = 0;
    int chunk = (int)blockIdx.x < num_static ? (int)blockIdx.x : -1;  // -1: draw a ticket
    bool chunk_open = false;   // `chunk` has been entered (its first segment was set up)
    bool finished = false;
    int scan_item = 0;         // forward scan over the requests (chunk starts only move forward)
    int64_t scan_pre = 0;
    Tile cur_t, nxt_t;
    int r[4], rn[4];
    int last_slot = -1, last_slot_n = -1;
    // describes the next tile into `nxt_t` (setting up its segment, and issuing the segment's query-block copy, when
    // it is the segment's first); returns false at the end of the range
    auto next_tile = [&]() -> bool {
      if (t == t_hi) {  // segment exhausted (or first call)
        if (finished) return false;
        if (t_hi != 0) {
          cur = seg_end;
          if (cur == ue && hg == HG - 1 && cur < g1) {
            pre += it.tiles;
            ++item;
            it = ws_item(p, item);
          }
        }
        bool first_of_chunk = false;
        if (!chunk_open || cur >= g1) {
          // ---- next chunk: the static one first, then tickets ----
          if (chunk_open || chunk < 0) {
            int tk = 0;
            if (lane == 0) tk = num_static + atomicAdd(sched_next, 1);
            chunk = __shfl_sync(0xffffffffu, tk, 0);
          }
          chunk_open = true;
          const int qs = nseg & 1;
          if (nseg >= 2) mbar_wait(bar(kWsQEmpty + qs), ((nseg >> 1) + 1) & 1);
          if (chunk >= num_chunks) {
            // no more work: tell the consumers, and let the last CTA to get here reset the scheduler
            if (lane == 0) {
              s_msg[qs].k = -1;
              mbar_arrive(bar(kWsQFull + qs));
              if (atomicAdd(sched_done, 1) == (int)gridDim.x - 1) {
                *sched_next = 0;
                *sched_done = 0;
              }
            }
            finished = true;
            return false;
          }
          chunk_bounds(chunk, cur, g1);
          // request holding tile `cur`: warp-parallel forward scan from where the previous chunk's search ended
          while (true) {
            const int i = scan_item + lane;
            const int ti = i < p.B ? ws_item(p, i).tiles : 0;
            int64_t incl = ti;
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) {
              const int64_t y = __shfl_up_sync(0xffffffffu, incl, o);
              if (lane >= o) incl += y;
            }
            const unsigned hit = __ballot_sync(0xffffffffu, i < p.B && (scan_pre + incl) * HG > cur);
            if (hit != 0u) {
              const int l = __ffs(hit) - 1;
              const int64_t incl_l = __shfl_sync(0xffffffffu, incl, l);
              const int ti_l = __shfl_sync(0xffffffffu, ti, l);
              scan_item += l;
              scan_pre += incl_l - ti_l;
              break;
            }
            scan_pre += __shfl_sync(0xffffffffu, incl, 31);
            scan_item += 32;
          }
          item = scan_item;
          pre = scan_pre;
          it = ws_item(p, item);
          first_of_chunk = true;
        }
        const int64_t ubase = pre * HG;
        hg = (int)((cur - ubase) / it.tiles);
        const int64_t us = ubase + (int64_t)hg * it.tiles;
        ue = us + it.tiles;
        t = (int)(cur - us);
        seg_end = min(g1, ue);
        t_hi = t + (int)(seg_end - cur);
        const int kvh = hg / p.groups_per_kvh;
        const int grp = hg - kvh * p.groups_per_kvh;
        const int qrow = p.q_rows ? p.q_rows[item] : item;
        const int h0 = kvh * p.G + grp * kDecMaxG;
        const int gcount = min(kDecMaxG, p.G - grp * kDecMaxG);
        const int qs = nseg & 1;
        if (!first_of_chunk && nseg >= 2) mbar_wait(bar(kWsQEmpty + qs), ((nseg >> 1) + 1) & 1);
        if (lane == 0) {
          if (first_of_chunk) {  // published with the segment's query block (the arrive below releases these writes)
            s_msg[qs].k = chunk;
            s_msg[qs].item = item;
            s_msg[qs].pre = pre;
          }
          const uint32_t qbytes = (uint32_t)(gcount * D * 2);
          mbar_arrive_expect_tx(bar(kWsQFull + qs), qbytes);
          bulk_load_1d(smem0 + S::kQ + qs * S::kQBytes,
                       reinterpret_cast<const T*>(p.q) + (int64_t)qrow * p.q_stride + (int64_t)h0 * D, qbytes,
                       bar(kWsQFull + qs));
        }
        ++nseg;
        nxt_t.idx = p.kv_indices + it.kv_beg;
        nxt_t.k_hi = min(it.L, t_hi * kDecTile);
        nxt_t.col = kvh * D;
        nxt_t.kvh = kvh;
        nxt_t.item = item;
        nxt_t.L = it.L;
      }
      nxt_t.key0 = t * kDecTile;
      nxt_t.last_of_unit = (t == it.tiles - 1);
      ++t;
      return true;
    };
    auto fetch_ids = [&](const Tile& td, int (&out)[4], int& lslot) {
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        const int key = td.key0 + (lane & 15) * 4 + i;
        out[i] = key < td.k_hi ? __ldg(td.idx + key) : num_slots;  // row past the pool -> zero-filled by the copy engine
      }
      lslot = (fused.new_k != nullptr && td.last_of_unit && td.L > 0) ? __ldg(td.idx + td.L - 1) : -1;
    };
    bool have = next_tile();
    if (have) {
      cur_t = nxt_t;
      fetch_ids(cur_t, r, last_slot);
    }
    while (have) {
      const bool have_next = next_tile();
      if (have_next) fetch_ids(nxt_t, rn, last_slot_n);
      // ---- issue the current tile ----
      if (last_slot >= 0 && last_slot != fused.skip_slot && last_slot < num_slots) {
        // new-token row of this unit: written right before the tile that contains it is fetched
        constexpr int kVecs = D * 2 / 16;  // 16-byte pieces of one (token, kv head) row
        const bool isv = lane >= 16;
        const uint8_t* src = reinterpret_cast<const uint8_t*>(isv ? fused.new_v : fused.new_k) +
                             ((int64_t)cur_t.item * fused.new_stride + (int64_t)cur_t.kvh * D) * 2;
        uint8_t* dstg = reinterpret_cast<uint8_t*>(const_cast<void*>(isv ? p.v_cache : p.k_cache)) +
                        ((int64_t)last_slot * p.kv_stride + (int64_t)cur_t.kvh * D) * 2;
        for (int v = lane & 15; v < kVecs; v += 16)
          *reinterpret_cast<uint4*>(dstg + v * 16) = *reinterpret_cast<const uint4*>(src + v * 16);
        // generic-proxy global writes -> ordered before the async-proxy (TMA) reads issued after the __syncwarp below
        asm volatile("fence.proxy.async.global;\n" ::: "memory");
      }
      const int g0s = __shfl_sync(0xffffffffu, r[0], lane & ~3);
      bool contig = (r[1] == r[0] + 1) && (r[2] == r[0] + 2) && (r[3] == r[0] + 3) &&
                    (r[0] == g0s + 4 * (lane & 3));
      contig = __all_sync(0xffffffffu, contig) && (cur_t.key0 + kDecTile <= cur_t.k_hi);
      const int st = n % kWsStages;
      if (n >= kWsStages) mbar_wait(bar(kWsEmpty + st), ((n / kWsStages) + 1) & 1);
      if (lane == 0) mbar_arrive_expect_tx(bar(kWsFull + st), S::kStageBytes);
      __syncwarp();
      const uint32_t dst = smem0 + S::kRing + st * S::kStageBytes;
      if (contig) {
        // 4 pages x halves x {K, V}: one 16-row box each, spread over the first lanes
        const int job = lane;  // job = op * 4 * kHalves + page * kHalves + hf
        const int op = job / (4 * kHalves), pg = (job / kHalves) & 3, hf = job % kHalves;
        const int row = __shfl_sync(0xffffffffu, r[0], pg * 4);
        if (job < 8 * kHalves)
          tma_load_2d(dst + op * S::kTileBytes + hf * S::kHalfBytes + pg * 16 * 128, op ? &tm_vb : &tm_kb,
                      bar(kWsFull + st), cur_t.col + hf * 64, row);
      } else {
        const int op = lane >> 4;
#pragma unroll
        for (int hf = 0; hf < kHalves; ++hf)
          tma_gather4(dst + op * S::kTileBytes + hf * S::kHalfBytes + (lane & 15) * 4 * 128, op ? &tm_vg : &tm_kg,
                      bar(kWsFull + st), cur_t.col + hf * 64, r[0], r[1], r[2], r[3]);
      }
      WS_STAMP(n == 0 && lane == 0, 3);
      ++n;
      have = have_next;
      cur_t = nxt_t;
#pragma unroll
      for (int i = 0; i < 4; ++i) r[i] = rn[i];
      last_slot = last_slot_n;
    }
    return;
  }

  if (warp == 5) {
    // ================================ merger ================================
    for (int mseq = 0;; ++mseq) {
      const int ms = mseq % kWsMq;
      mbar_wait(bar(kWsMqFull + ms), (mseq / kWsMq) & 1);
      const WsMerge d = s_mq[ms];
      __syncwarp();
      if (lane == 0) mbar_arrive(bar(kWsMqEmpty + ms));
      if (d.unit < 0) break;
      // the piece's partial rows (written by the consumer warps, ordered before the hand-off) become visible
      // device-wide before the ticket is drawn
      __threadfence();
      int last = 0;
      if (lane == 0) {
        int* cnt = p.ws_cnt + d.unit;
        const int ticket = atomicAdd(cnt, 1);
        last = (ticket == d.c_last - d.c_first);
        if (last) *cnt = 0;  // self-cleaning
      }
      last = __shfl_sync(0xffffffffu, last, 0);
      if (!last) continue;
      __threadfence();
      for (int e = lane; e < d.gcount * (D / 4); e += 32) {
        const int r = e / (D / 4), cc = (e - r * (D / 4)) * 4;
        const int h = d.h0 + r;
        float m = -INFINITY;
        for (int pc = d.c_first; pc <= d.c_last; ++pc) {
          const int64_t w = (int64_t)(pc * 2 + (pc == d.c_first ? 1 : 0)) * ws_rows + r;
          m = fmaxf(m, __ldcg(p.ws_lse + w));
        }
        float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
        float l = 0.f;
        for (int pc = d.c_first; pc <= d.c_last; ++pc) {
          const int64_t w = (int64_t)(pc * 2 + (pc == d.c_first ? 1 : 0)) * ws_rows + r;
          const float ls = __ldcg(p.ws_lse + w);
          const float f = (ls == -INFINITY) ? 0.f : fast_exp2(ls - m);
          const float4 x = __ldcg(reinterpret_cast<const float4*>(p.ws_o + w * D + cc));
          acc.x += f * x.x;
          acc.y += f * x.y;
          acc.z += f * x.z;
          acc.w += f * x.w;
          l += f;
        }
        if (p.sinks != nullptr && l > 0.f) l += fast_exp2(__ldg(p.sinks + h) * kLog2e - m);
        const float inv = (l > 0.f) ? 1.f / l : 0.f;
        T* dst = reinterpret_cast<T*>(p.o) + (int64_t)d.qrow * p.o_stride + (int64_t)h * D + cc;
        uint2 pk;
        pk.x = pack2<kBf16>(acc.x * inv, acc.y * inv);
        pk.y = pack2<kBf16>(acc.z * inv, acc.w * inv);
        *reinterpret_cast<uint2*>(dst) = pk;
        if (WS_LSE_ON && cc == 0)
          p.lse[(int64_t)d.qrow * p.Hq + h] = (l > 0.f) ? (m + log2f(l)) * kLn2 : -INFINITY;
      }
    }
    return;
  }

  // ================================ consumers ================================
  int mseq = 0;  // pieces handed to the merger warp so far
  const int lm = lane >> 3, lr = lane & 7;
  float* sm_m = reinterpret_cast<float*>(smem + S::kML);  // [4][16]
  float* sm_l = sm_m + 64;                                // [4][16]
  int chunk = -1;
  while (true) {
    // every segment starts with the producer's query block; the first segment of a chunk also carries the chunk
    {
      const int qs = nseg & 1;
      mbar_wait(bar(kWsQFull + qs), (nseg >> 1) & 1);
      if (cur >= g1) {
        const WsMsg msg = s_msg[qs];
        if (msg.k < 0) break;
        chunk = msg.k;
        item = msg.item;
        pre = msg.pre;
        chunk_bounds(chunk, cur, g1);
        it = ws_item(p, item);
      }
    }
    const int64_t ubase = pre * HG;
    const int hg = (int)((cur - ubase) / it.tiles);
    const int64_t us = ubase + (int64_t)hg * it.tiles;
    const int64_t ue = us + it.tiles;
    const int t_lo = (int)(cur - us);
    const int64_t seg_end = min(g1, ue);
    const int t_hi = t_lo + (int)(seg_end - cur);
    const bool whole = (t_lo == 0 && t_hi == it.tiles);
    const int kvh = hg / p.groups_per_kvh;
    const int grp = hg - kvh * p.groups_per_kvh;
    const int qrow = p.q_rows ? p.q_rows[item] : item;
    const int h0 = kvh * p.G + grp * kDecMaxG;
    const int gcount = min(kDecMaxG, p.G - grp * kDecMaxG);
    const int k_lo = t_lo * kDecTile;
    const int k_hi = min(it.L, t_hi * kDecTile);

    // ---- Q fragments (A operand, 16 x D) from the producer's bulk copy; rows >= gcount are zero ----------
    uint32_t qf[D / 16][4];
    {
      const int qs = nseg & 1;
      const uint8_t* qb = smem + S::kQ + qs * S::kQBytes;
      const int r0 = lane >> 2, c0 = (lane & 3) * 2;
#pragma unroll
      for (int ks = 0; ks < D / 16; ++ks) {
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          const int r = r0 + (j & 1) * 8;
          const int cc = ks * 16 + c0 + (j >> 1) * 8;
          qf[ks][j] = (r < gcount) ? *reinterpret_cast<const uint32_t*>(qb + (r * D + cc) * 2) : 0u;
        }
      }
      __syncwarp();
      if (lane == 0) mbar_arrive(bar(kWsQEmpty + qs));
    }

    float o_acc[D / 8][4];
#pragma unroll
    for (int i = 0; i < D / 8; ++i) o_acc[i][0] = o_acc[i][1] = o_acc[i][2] = o_acc[i][3] = 0.f;
    float m_run[2] = {-INFINITY, -INFINITY};
    float l_run[2] = {0.f, 0.f};
    int st = 0;
    for (int t = t_lo; t < t_hi; ++t, ++n) {
      st = n % kWsStages;
      mbar_wait(bar(kWsFull + st), (n / kWsStages) & 1);
      WS_STAMP(n == 0 && tid == 0, 4);
      const uint32_t ks = smem0 + S::kRing + st * S::kStageBytes;
      const uint32_t vs = ks + S::kTileBytes;
      const int key0 = t * kDecTile + warp * 16;  // first key of this warp
      if (key0 < k_hi) {                          // warp-uniform
        float s[2][4] = {{0.f, 0.f, 0.f, 0.f}, {0.f, 0.f, 0.f, 0.f}};
        {
          const int krow = warp * 16 + (lm >> 1) * 8 + lr;
#pragma unroll
          for (int ksp = 0; ksp < D / 16; ++ksp) {
            uint32_t b0, b1, b2, b3;
            ldmatrix_x4(b0, b1, b2, b3, ks + ws_off<D>(krow, 2 * ksp + (lm & 1)));
            mma_16816<kBf16>(s[0], qf[ksp], b0, b1);
            mma_16816<kBf16>(s[1], qf[ksp], b2, b3);
          }
        }
        const int kcol = key0 + (lane & 3) * 2;
        float tmax[2] = {-INFINITY, -INFINITY};
#pragma unroll
        for (int nb = 0; nb < 2; ++nb) {
#pragma unroll
          for (int j = 0; j < 4; ++j) {
            const int key = kcol + nb * 8 + (j & 1);
            float x = p.cap_post > 0.f ? p.cap_post * tanh_approx(s[nb][j] * p.cap_pre) : s[nb][j] * p.scale_log2;
            x = (key < k_hi) ? x : -INFINITY;
            s[nb][j] = x;
            tmax[j >> 1] = fmaxf(tmax[j >> 1], x);
          }
        }
        float alpha[2];
#pragma unroll
        for (int r = 0; r < 2; ++r) {
          tmax[r] = fmaxf(tmax[r], __shfl_xor_sync(0xffffffffu, tmax[r], 1));
          tmax[r] = fmaxf(tmax[r], __shfl_xor_sync(0xffffffffu, tmax[r], 2));
          const float m_new = fmaxf(m_run[r], tmax[r]);  // finite: key0 < k_hi => >= 1 valid key
          alpha[r] = fast_exp2(m_run[r] - m_new);
          m_run[r] = m_new;
          l_run[r] *= alpha[r];
        }
        uint32_t pf[4];
#pragma unroll
        for (int nb = 0; nb < 2; ++nb) {
          const float p0 = fast_exp2(s[nb][0] - m_run[0]);
          const float p1 = fast_exp2(s[nb][1] - m_run[0]);
          const float p2 = fast_exp2(s[nb][2] - m_run[1]);
          const float p3 = fast_exp2(s[nb][3] - m_run[1]);
          l_run[0] += p0 + p1;
          l_run[1] += p2 + p3;
          pf[nb * 2 + 0] = pack2<kBf16>(p0, p1);
          pf[nb * 2 + 1] = pack2<kBf16>(p2, p3);
        }
        const int vrow = warp * 16 + (lm & 1) * 8 + lr;
        const bool rescale = __any_sync(0xffffffffu, alpha[0] != 1.f || alpha[1] != 1.f);
#pragma unroll
        for (int np = 0; np < D / 16; ++np) {
          uint32_t b0, b1, b2, b3;
          ldmatrix_x4_trans(b0, b1, b2, b3, vs + ws_off<D>(vrow, 2 * np + (lm >> 1)));
          if (rescale) {  // warp-uniform: the running max moved in this tile
#pragma unroll
            for (int h = 0; h < 2; ++h) {
              float(&acc)[4] = o_acc[2 * np + h];
              acc[0] *= alpha[0];
              acc[1] *= alpha[0];
              acc[2] *= alpha[1];
              acc[3] *= alpha[1];
            }
          }
          mma_16816<kBf16>(o_acc[2 * np], pf, b0, b1);
          mma_16816<kBf16>(o_acc[2 * np + 1], pf, b2, b3);
        }
      }
      if (t + 1 < t_hi) {  // the segment's last stage is released after the merge below
        __syncwarp();
        if (lane == 0) mbar_arrive(bar(kWsEmpty + st));
      }
    }

    WS_STAMP(tid == 0, 5);
    // ---- merge the four warps through the last tile's stage (fp32 [4][16][D] == one stage) ------------------
#pragma unroll
    for (int r = 0; r < 2; ++r) {
      l_run[r] += __shfl_xor_sync(0xffffffffu, l_run[r], 1);
      l_run[r] += __shfl_xor_sync(0xffffffffu, l_run[r], 2);
    }
    float* sm_o = reinterpret_cast<float*>(smem + S::kRing + st * S::kStageBytes);
    consumer_sync();  // every warp has finished reading the stage
    {
      const int r0 = lane >> 2, c0 = (lane & 3) * 2;
#pragma unroll
      for (int i = 0; i < D / 8; ++i) {
        float* dst = sm_o + (warp * 16 + r0) * D + i * 8 + c0;
        *reinterpret_cast<float2*>(dst) = make_float2(o_acc[i][0], o_acc[i][1]);
        *reinterpret_cast<float2*>(dst + 8 * D) = make_float2(o_acc[i][2], o_acc[i][3]);
      }
      if ((lane & 3) == 0) {
        sm_m[warp * 16 + r0] = m_run[0];
        sm_m[warp * 16 + r0 + 8] = m_run[1];
        sm_l[warp * 16 + r0] = l_run[0];
        sm_l[warp * 16 + r0 + 8] = l_run[1];
      }
    }
    consumer_sync();
    const int wslot = chunk * 2 + (t_lo == 0 ? 1 : 0);  // partial pieces only
    for (int e = tid; e < gcount * (D / 4); e += kWsConsumers) {
      const int r = e / (D / 4), cc = (e - r * (D / 4)) * 4;
      float m = -INFINITY;
#pragma unroll
      for (int w = 0; w < 4; ++w) m = fmaxf(m, sm_m[w * 16 + r]);
      float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
      float l = 0.f;
#pragma unroll
      for (int w = 0; w < 4; ++w) {
        const float mw = sm_m[w * 16 + r];
        const float f = (mw == -INFINITY) ? 0.f : fast_exp2(mw - m);
        const float4 x = *reinterpret_cast<const float4*>(sm_o + (w * 16 + r) * D + cc);
        acc.x += f * x.x;
        acc.y += f * x.y;
        acc.z += f * x.z;
        acc.w += f * x.w;
        l += f * sm_l[w * 16 + r];
      }
      if (whole && p.sinks != nullptr && l > 0.f) l += fast_exp2(__ldg(p.sinks + h0 + r) * kLog2e - m);
      const float inv = (l > 0.f) ? 1.f / l : 0.f;
      acc.x *= inv;
      acc.y *= inv;
      acc.z *= inv;
      acc.w *= inv;
      const float lse2 = (l > 0.f) ? m + log2f(l) : -INFINITY;
      const int h = h0 + r;
      if (whole) {
        T* dst = reinterpret_cast<T*>(p.o) + (int64_t)qrow * p.o_stride + (int64_t)h * D + cc;
        uint2 pk;
        pk.x = pack2<kBf16>(acc.x, acc.y);
        pk.y = pack2<kBf16>(acc.z, acc.w);
        *reinterpret_cast<uint2*>(dst) = pk;
        if (WS_LSE_ON && cc == 0) p.lse[(int64_t)qrow * p.Hq + h] = lse2 * kLn2;
      } else {
        const int64_t w = (int64_t)wslot * ws_rows + r;
        *reinterpret_cast<float4*>(p.ws_o + w * D + cc) = acc;
        if (cc == 0) p.ws_lse[w] = lse2;
      }
    }
    if (!whole) {
      // ---- hand the piece to the merger warp: it takes the unit's arrival ticket and, if this was the last piece,
      //      merges them - a fence, an atomic round trip and an L2 read that the streaming warps do not wait for ----
      consumer_sync();  // every partial row of this piece has been written (by all four warps)
      if (tid == 0) {
        const int ms = mseq % kWsMq;
        if (mseq >= kWsMq) mbar_wait(bar(kWsMqEmpty + ms), ((mseq / kWsMq) + 1) & 1);
        WsMerge& d = s_mq[ms];
        d.unit = item * HG + hg;
        d.qrow = qrow;
        d.h0 = h0;
        d.gcount = gcount;
        d.c_first = chunk_of(us);
        d.c_last = chunk_of(ue - 1);
        mbar_arrive(bar(kWsMqFull + ms));
      }
      ++mseq;
    }
    consumer_sync();  // scratch (the stage), sm_m / sm_l and s_last are free again
    fence_proxy_async_smem();  // generic-proxy scratch writes before the copy engine refills the stage
    if (lane == 0) mbar_arrive(bar(kWsEmpty + st));
    ++nseg;
    cur = seg_end;
    if (cur == ue && hg == HG - 1 && cur < g1) {
      pre += it.tiles;
      ++item;
      it = ws_item(p, item);
    }
  }
  if (tid == 0) {  // tell the merger warp that no more pieces will come
    const int ms = mseq % kWsMq;
    if (mseq >= kWsMq) mbar_wait(bar(kWsMqEmpty + ms), ((mseq / kWsMq) + 1) & 1);
    s_mq[ms].unit = -1;
    mbar_arrive(bar(kWsMqFull + ms));
  }
  WS_STAMP(tid == 0, 6);
}

int ws_ctas() { return 2 * sm_count(); }

template <int D, bool kBf16>
static int launch_decode_ws(const CUtensorMap& kg, const CUtensorMap& vg, const CUtensorMap& kb, const CUtensorMap& vb,
                            const DecodeParams& p, const WsFused& fused, int min_tiles, int num_slots, int ws_rows,
                            cudaStream_t st) {
  constexpr int kSmem = WsSmem<D>::kTotal + 1024;  // + alignment slack
  static bool configured = false;
  if (!configured) {
    B2_CUDA(cudaFuncSetAttribute(decode_ws_kernel<D, kBf16>, cudaFuncAttributeMaxDynamicSharedMemorySize, kSmem));
    B2_CUDA(cudaFuncSetAttribute(decode_ws_kernel<D, kBf16>, cudaFuncAttributePreferredSharedMemoryCarveout, 100));
    configured = true;
  }
  // B200ATTN_PDL=1: programmatic stream serialization - the grid may start while the previous kernel of the stream
  // drains; the kernel orders itself with griddepcontrol.wait (dev switch, measured in profiles/)
  static const bool pdl = [] { const char* e = getenv("B200ATTN_PDL"); return e != nullptr && atoi(e) != 0; }();
  if (pdl) {
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3((unsigned)ws_ctas());
    cfg.blockDim = dim3(kWsThreads);
    cfg.dynamicSmemBytes = kSmem;
    cfg.stream = st;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr;
    cfg.numAttrs = 1;
    B2_CUDA(cudaLaunchKernelEx(&cfg, decode_ws_kernel<D, kBf16>, kg, vg, kb, vb, p, fused, min_tiles, num_slots, ws_rows));
    return B200ATTN_OK;
  }
  decode_ws_kernel<D, kBf16><<<ws_ctas(), kWsThreads, kSmem, st>>>(kg, vg, kb, vb, p, fused, min_tiles, num_slots, ws_rows);
  B2_LAUNCH_CHECK();
  return B200ATTN_OK;
}

}  // namespace b200attn

namespace b200attn {
static int ws_rows_for(int Hq, int Hkv) { return Hq / Hkv < kDecMaxG ? Hq / Hkv : kDecMaxG; }
static int64_t ws_max_chunks() { return (int64_t)(kWsMaxLevels + 1) * ws_ctas(); }
}  // namespace b200attn

extern "C" int64_t b200attn_decode_ws_workspace_bytes(int32_t Hq, int32_t Hkv, int32_t D) {
  using namespace b200attn;
  if (Hq <= 0 || Hkv <= 0 || D <= 0) return 0;
  // fixed counter header + two partial slots per chunk of the guided schedule ([min(G,16)][D] fp32 + lse)
  return kDecCounterBytes + ws_max_chunks() * 2 * ws_rows_for(Hq, Hkv) * (D + 1) * 4;
}

extern "C" int b200attn_decode_ws(const void* q, int64_t q_stride, void* k_cache, void* v_cache, int64_t kv_stride,
                                  int64_t num_slots, void* o, int64_t o_stride, float* lse,
                                  const int32_t* kv_indptr, const int32_t* kv_indices, const int32_t* q_rows,
                                  const int32_t* req_ids, int32_t B, int32_t Hq, int32_t Hkv, int32_t D, int dtype,
                                  float sm_scale, int32_t min_chunk, int32_t window_left, float logit_cap,
                                  const float* sinks, const void* new_k, const void* new_v, int64_t new_kv_stride,
                                  int64_t skip_slot, void* workspace, int64_t workspace_bytes, void* stream) {
  using namespace b200attn;
  B2_CHECK_ARG(B >= 0, "decode_ws: negative batch");
  B2_CHECK_ARG(logit_cap >= 0.f, "decode_ws: negative logit_cap");
  if (B == 0) return B200ATTN_OK;
  B2_CHECK_ARG(q && k_cache && v_cache && o && kv_indptr && kv_indices, "decode_ws: null pointer");
  B2_CHECK_ARG(Hq > 0 && Hkv > 0 && Hq % Hkv == 0, "decode_ws: Hq=%d must be a multiple of Hkv=%d", Hq, Hkv);
  B2_CHECK_ARG(min_chunk >= 1, "decode_ws: bad min_chunk");
  B2_CHECK_ARG(num_slots > 0 && num_slots < (1ll << 31), "decode_ws: bad num_slots");
  B2_CHECK_ARG(q_stride % 8 == 0 && kv_stride % 8 == 0 && o_stride % 4 == 0,
               "decode_ws: strides must keep rows 16-byte aligned");
  B2_CHECK_ARG((((uintptr_t)q | (uintptr_t)k_cache | (uintptr_t)v_cache | (uintptr_t)o) & 15) == 0,
               "decode_ws: tensors must be 16-byte aligned");
  B2_CHECK_ARG((new_k == nullptr) == (new_v == nullptr), "decode_ws: new_k and new_v must be given together");
  B2_CHECK_ARG(new_k == nullptr || (q_rows == nullptr && req_ids == nullptr && new_kv_stride % 8 == 0 &&
                                    (((uintptr_t)new_k | (uintptr_t)new_v) & 15) == 0),
               "decode_ws: the fused store needs identity row maps and 16-byte aligned new_k / new_v rows");
  if (dtype != B200ATTN_BF16 && dtype != B200ATTN_FP16) B2_UNSUPPORTED("decode_ws: dtype %d", dtype);
  if (D != 64 && D != 128) B2_UNSUPPORTED("decode_ws: head_dim %d (supported: 64, 128)", D);
  const int64_t need = b200attn_decode_ws_workspace_bytes(Hq, Hkv, D);
  B2_CHECK_ARG(workspace && workspace_bytes >= need, "decode_ws: workspace too small (%lld < %lld)",
               (long long)workspace_bytes, (long long)need);
  B2_CHECK_ARG((int64_t)B * Hkv * ((Hq / Hkv + kDecMaxG - 1) / kDecMaxG) <= kDecMaxUnits - 2,
               "decode_ws: more than %d units", kDecMaxUnits - 2);  // the last two counters drive the chunk tickets
  const CUtensorMap *kg, *vg, *kb, *vb;
  int rc = tensor_map_cached(&kg, k_cache, (uint64_t)num_slots, (uint64_t)Hkv * D, (uint64_t)kv_stride * 2, 64, 1);
  if (rc) return rc;
  rc = tensor_map_cached(&vg, v_cache, (uint64_t)num_slots, (uint64_t)Hkv * D, (uint64_t)kv_stride * 2, 64, 1);
  if (rc) return rc;
  rc = tensor_map_cached(&kb, k_cache, (uint64_t)num_slots, (uint64_t)Hkv * D, (uint64_t)kv_stride * 2, 64, 16);
  if (rc) return rc;
  rc = tensor_map_cached(&vb, v_cache, (uint64_t)num_slots, (uint64_t)Hkv * D, (uint64_t)kv_stride * 2, 64, 16);
  if (rc) return rc;
  const int rows = ws_rows_for(Hq, Hkv);
  DecodeParams p;
  p.q = q;
  p.k_cache = k_cache;
  p.v_cache = v_cache;
  p.o = o;
  p.lse = lse;
  p.kv_indptr = kv_indptr;
  p.kv_indices = kv_indices;
  p.q_rows = q_rows;
  p.req_ids = req_ids;
  p.ws_cnt = reinterpret_cast<int*>(workspace);
  p.ws_o = reinterpret_cast<float*>(reinterpret_cast<uint8_t*>(workspace) + kDecCounterBytes);
  p.ws_lse = p.ws_o + ws_max_chunks() * 2 * rows * D;
  p.q_stride = q_stride;
  p.kv_stride = kv_stride;
  p.o_stride = o_stride;
  p.B = B;
  p.Hq = Hq;
  p.Hkv = Hkv;
  p.G = Hq / Hkv;
  p.groups_per_kvh = (p.G + kDecMaxG - 1) / kDecMaxG;
  p.max_splits = 1;
  p.min_chunk = min_chunk;
  p.scale_log2 = sm_scale * kLog2e;
  p.window_left = window_left;
  p.cap_pre = logit_cap > 0.f ? sm_scale / logit_cap : 0.f;
  p.cap_post = logit_cap > 0.f ? logit_cap * kLog2e : 0.f;
  p.sinks = sinks;
  p.tail_begin = -1;
  p.out_scale = 1.f;
  WsFused fused;
  fused.new_k = new_k;
  fused.new_v = new_v;
  fused.new_stride = new_kv_stride;
  fused.skip_slot = (int)skip_slot;
  const int min_tiles = (min_chunk + kDecTile - 1) / kDecTile;
  cudaStream_t st = (cudaStream_t)stream;
  if (D == 128)
    return dtype == B200ATTN_BF16
               ? launch_decode_ws<128, true>(*kg, *vg, *kb, *vb, p, fused, min_tiles, (int)num_slots, rows, st)
               : launch_decode_ws<128, false>(*kg, *vg, *kb, *vb, p, fused, min_tiles, (int)num_slots, rows, st);
  return dtype == B200ATTN_BF16
             ? launch_decode_ws<64, true>(*kg, *vg, *kb, *vb, p, fused, min_tiles, (int)num_slots, rows, st)
             : launch_decode_ws<64, false>(*kg, *vg, *kb, *vb, p, fused, min_tiles, (int)num_slots, rows, st);
}
