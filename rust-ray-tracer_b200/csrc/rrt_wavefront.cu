// rrt_wavefront.cu — Scene::draw_scene (engine.rs:186-255) as a WAVEFRONT of small kernels over compacted queues.
//
// Why not one persistent kernel (the round-1 design, still used for the per-ray seams): with the searches of the
// scene-wide hierarchy in place a frame alternates between four pieces of code — camera search, shading, shadow
// filter, exact walk — and a warp that does all four in turn (a) refetches ~70 KB of instructions per ray (ncu: 29 % of
// the stall samples were "no instruction"), (b) keeps its lanes idle whenever their sample has no ray of the current
// kind (c4: 46 % of the camera rays hit nothing, so shadow queries ran at 17 of 32 lanes), and (c) carries the whole
// shading state (~700 bytes of local memory per lane) through the traversal loops at 64 registers.  Here every kind
// of work is a kernel of its own with its own register budget, every warp is full (queues of sample slots are
// compacted per block in shared memory), and what must survive between two kernels is ~190 bytes per LIVE sample in
// slot-indexed arrays in HBM (c4: ~1.5 ms worth of traffic per frame at the measured 6.45 TB/s, overlapped).
//
// A frame:  camera_kernel (ray generation + sparse_closest; samples that hit nothing finish here and never touch the
// state arrays) -> walk_kernel (the few rays the search handed over) -> rounds of { shade_kernel (RayState::advance:
// hit set-up, lights, next shadow / mirror ray, or the finished colour into the per-pixel accumulator),
// filter_kernel (scene_clear), walk_kernel } until no sample is left.  The arithmetic of every ray is the code of
// rrt_trace.cuh / rrt_shade.cuh, unchanged: results are bit-identical to the per-ray seams and to the oracle.
#include "rrt_render.h"
#include "rrt_raygen.cuh"

namespace rrt {

namespace {

constexpr int kWfBlock = 128;
constexpr int kSlotsPerBlock = 1024;        // contiguous slots one block looks after in the queue-less kernels
// What a slot is waiting for, in the top bits of its flags word.  There are no global queues: every kernel lets each
// block scan the flags of ITS 1024 consecutive slots and compact the ones it is responsible for into shared memory
// (one shared-memory atomic per warp), so warps are full and nothing contends for a global counter — a first version
// with three global queues spent more time in same-address atomics (~2.5 ns each, ~6 M per frame) than in tracing.
constexpr uint32_t kWorkMask = 7u << 28;
constexpr uint32_t kDone = 0u << 28;        // finished (or never traced)
constexpr uint32_t kShade = 1u << 28;       // a traversal result is waiting for RayState::advance
constexpr uint32_t kFilter = 2u << 28;      // shadow ray stored, to be tried by scene_clear
constexpr uint32_t kWalk = 3u << 28;        // closest-hit ray stored, exact walk
constexpr uint32_t kWalkAny = 4u << 28;     // shadow ray stored, exact walk
constexpr uint32_t kWalkCam = 5u << 28;     // camera ray (recomputed from the slot, nothing stored yet), exact walk
constexpr uint32_t kFresh = 1u << 24;       // the slot's camera ray has just been traced, no shading state stored yet

// Device memory of one frame in flight, indexed by SLOT = position in the sample enumeration of rrt_raygen.cuh
// (16x4-pixel blocks of eight 4x2-pixel tiles x 4 supersamples), all arrays structure-of-arrays.
struct WfState {
    uint32_t n;            // slots
    double* ray;           // [7][n]  origin, direction, max_t of the ray to trace next (shadow / mirror rays)
    double* hit_t;         // [n]     result of the traversal just made
    uint32_t* hit_id;      // [n]
    uint32_t* hit_rec;     // [n]
    double* parked;        // [12][n] RayState::parked
    uint32_t* texel;       // [n]
    uint32_t* res;         // [n]
    int32_t* hit0;         // [n]
    int32_t* mid;          // [n]
    uint32_t* flags;       // [n]     depth | li << 8 | shadow << 16 | fresh << 24 | work << 28
    double* mirror;        // [4 * levels][n] open mirror levels (lx, ly, lz, lk per level)
    uint32_t* pending;     // [64]    pending[r] != 0: the shading of round r sent more rays out
};

// The slots of this block whose work state is in [lo, hi], compacted into `list` (shared, kSlotsPerBlock words).
// Returns their number.  Called by all threads of the block.
__device__ __forceinline__ uint32_t compact_block(const uint32_t* __restrict__ flags, uint32_t total, uint32_t lo, uint32_t hi,
                                                  uint32_t* list, uint32_t* n_list) {
    if (threadIdx.x == 0) *n_list = 0u;
    __syncthreads();
    const uint32_t base = blockIdx.x * kSlotsPerBlock;
    const unsigned lane = threadIdx.x & 31u;
#pragma unroll 1
    for (int it = 0; it < kSlotsPerBlock / kWfBlock; ++it) {
        const uint32_t slot = base + it * kWfBlock + threadIdx.x;
        bool want = false;
        if (slot < total) {
            const uint32_t w = flags[slot] & kWorkMask;
            want = w >= lo && w <= hi;
        }
        const unsigned m = __ballot_sync(0xFFFFFFFFu, want);
        if (m != 0u) {
            uint32_t off = 0;
            if (lane == (unsigned)__ffs(m) - 1u) off = atomicAdd(n_list, (uint32_t)__popc(m));
            off = __shfl_sync(0xFFFFFFFFu, off, __ffs(m) - 1);
            if (want) list[off + __popc(m & ((1u << lane) - 1u))] = slot;
        }
    }
    __syncthreads();
    return *n_list;
}

// slot -> pixel of the shard (compact index), frame index the finished pixel is stored at, supersample, camera ray
struct SlotPixel {
    size_t pix, out;
    unsigned k;
    bool in_frame, traced;
    V3 d;
};
__device__ __forceinline__ SlotPixel slot_pixel(const DevParams& P, uint32_t slot, int tiles_x) {
    SlotPixel sp;
    const uint32_t tile = slot >> 5;
    int col, krow;
    sample_pixel(tile >> 3, tile & 7u, slot & 31u, tiles_x, col, krow, sp.k);
    sp.in_frame = col < P.W && krow < P.n_rows_local;
    sp.traced = false;
    sp.pix = sp.out = 0;
    if (sp.in_frame) {
        sp.pix = (size_t)krow * P.W + col;
        const int r = buffer_row(P, krow);
        sp.out = P.full_frame ? (size_t)r * P.W + col : sp.pix;
        sp.traced = primary_ray(P, col, r, sp.k, sp.d);
    }
    return sp;
}

// one finished supersample: packed per-pixel accumulator, the 4th arrival applies Color::mix and stores the pixel
// (entities.rs:32-36,49-69; engine.rs:146-158)
__device__ __forceinline__ void finish_sample(uint32_t* __restrict__ acc, uint32_t* __restrict__ fb, const SlotPixel& sp, uint32_t rgb) {
    const uint32_t add = sample_add(rgb);
    const uint32_t old = atomicAdd(acc + sp.pix, add);
    if ((old >> 30) == 3u) fb[sp.out] = mix_pack(old + add);
}

// is the search whose switch sits in fstat[slot] (2: shadow filter, 6: sparse camera rays) to be tried this frame?
__device__ __forceinline__ bool search_on(const DevParams& P, int slot) {
    return P.search_mode == 2 || ((P.search_mode == 1 || P.search_mode == 3) && P.fstat[slot] != 0u);
}

struct WfTally {   // rrt_render_counters: summed per warp, one atomic per kernel exit
    unsigned long long primary = 0, shadow = 0, reflect = 0;
    TraceCount tc;
};
template <bool COUNT>
__device__ __forceinline__ void flush_tally(const WfTally& w, DevCounters* __restrict__ counters) {
    if (!COUNT) return;
    unsigned long long v[7] = {w.primary, w.shadow, w.reflect, w.tc.mt, w.tc.box, w.tc.cull, w.tc.nodes};
#pragma unroll
    for (int i = 0; i < 7; ++i) {
        for (int o = 16; o > 0; o >>= 1) v[i] += __shfl_xor_sync(0xFFFFFFFFu, v[i], o);
        if ((threadIdx.x & 31u) == 0u && v[i]) atomicAdd(reinterpret_cast<unsigned long long*>(counters) + i, v[i]);
    }
}
// what a search did with its rays, per warp -> fstat[at], fstat[at + 1]
__device__ __forceinline__ void flush_search_stats(uint32_t* __restrict__ fstat, int at, uint32_t answered, uint32_t handed) {
    if (!fstat) return;
    for (int o = 16; o > 0; o >>= 1) {
        answered += __shfl_xor_sync(0xFFFFFFFFu, answered, o);
        handed += __shfl_xor_sync(0xFFFFFFFFu, handed, o);
    }
    if ((threadIdx.x & 31u) == 0u && answered) atomicAdd(fstat + at, answered);
    if ((threadIdx.x & 31u) == 0u && handed) atomicAdd(fstat + at + 1, handed);
}

// ---------------------------------------------------------------- camera rays
// engine.rs:205-236 for every slot of the shard + the search that answers most camera rays (rrt_trace.cuh).
template <bool COUNT>
__global__ void __launch_bounds__(kWfBlock)
camera_kernel(const __grid_constant__ DevScene sc, const __grid_constant__ DevParams P, const __grid_constant__ WfState ws,
              uint32_t* __restrict__ fb, int32_t* __restrict__ hit_ids, uint32_t* __restrict__ acc, const int tiles_x,
              const uint32_t total, uint32_t* __restrict__ fstat, DevCounters* __restrict__ counters) {
    const bool sparse_on = search_on(P, 6) && sc.global_root != 0u && sc.sparse_ok != 0u;
    uint32_t answered = 0, handed = 0;
    WfTally tally;
    const uint32_t slot = blockIdx.x * kWfBlock + threadIdx.x;
    if (slot < total) {
        uint32_t state = kDone;
        const SlotPixel sp = slot_pixel(P, slot, tiles_x);
        if (sp.in_frame && !sp.traced) {
            // never written by the reference (row 0, odd last column / row): stays 0, hit id -1
            if (hit_ids) hit_ids[sp.pix * 4 + sp.k] = -1;
            if ((atomicAdd(acc + sp.pix, 1u << 30) >> 30) == 3u) fb[sp.out] = 0u;
        } else if (sp.traced) {
            if (COUNT) tally.primary++;
            SparseResult r;
            r.status = 0u;
            if (sparse_on) {
                r = sparse_closest<COUNT>(search_scene(sc), V3{P.ox, P.oy, P.oz}, sp.d, tally.tc);
                answered += r.status;
                handed += 1u - r.status;
            }
            if (r.status && r.id == kNone) {   // WHITE (raytracer.rs:109-111): done, no state ever stored
                if (hit_ids) hit_ids[sp.pix * 4 + sp.k] = -1;
                finish_sample(acc, fb, sp, 0x00FFFFFFu);
            } else if (r.status) {
                ws.hit_t[slot] = r.t; ws.hit_id[slot] = r.id; ws.hit_rec[slot] = r.rec;
                state = kShade | kFresh;
            } else {
                state = kWalkCam;
            }
        }
        ws.flags[slot] = state;
    }
    flush_search_stats(fstat, 4, answered, handed);
    flush_tally<COUNT>(tally, counters);
}

// ---------------------------------------------------------------- the exact walk for whatever was not answered
template <bool COUNT>
__global__ void __launch_bounds__(kWfBlock)
walk_kernel(const __grid_constant__ DevScene sc, const __grid_constant__ DevParams P, const __grid_constant__ WfState ws,
            const int tiles_x, const uint32_t total, DevCounters* __restrict__ counters) {
    __shared__ uint32_t list[kSlotsPerBlock];
    __shared__ uint32_t n_list;
    const uint32_t cnt = compact_block(ws.flags, total, kWalk, kWalkCam, list, &n_list);
    WfTally tally;
    for (uint32_t i = threadIdx.x; i < cnt; i += kWfBlock) {
        const uint32_t slot = list[i];
        const uint32_t fl = ws.flags[slot];
        const uint32_t work = fl & kWorkMask;
        V3 o, d;
        double max_t;
        uint32_t keep = fl & ~(kWorkMask | kFresh);
        if (work == kWalkCam) {
            const SlotPixel sp = slot_pixel(P, slot, tiles_x);
            o = V3{P.ox, P.oy, P.oz}; d = sp.d;
            max_t = __longlong_as_double(0x7FF0000000000000ll);
            keep = kFresh;
        } else {
            const uint32_t n = ws.n;
            o = V3{ws.ray[slot], ws.ray[n + slot], ws.ray[2 * n + slot]};
            d = V3{ws.ray[3 * n + slot], ws.ray[4 * n + slot], ws.ray[5 * n + slot]};
            max_t = ws.ray[6 * n + slot];
        }
        Hit h;
        walk<COUNT>(sc, o, d, max_t, work == kWalkAny, h, tally.tc);
        ws.hit_t[slot] = h.t; ws.hit_id[slot] = h.id; ws.hit_rec[slot] = h.rec;
        ws.flags[slot] = keep | kShade;
    }
    flush_tally<COUNT>(tally, counters);
}

// ---------------------------------------------------------------- shadow queries: scene_clear, else the walk
template <bool COUNT>
__global__ void __launch_bounds__(kWfBlock)
filter_kernel(const __grid_constant__ DevScene sc, const __grid_constant__ WfState ws, const uint32_t total,
              uint32_t* __restrict__ fstat, DevCounters* __restrict__ counters) {
    __shared__ uint32_t list[kSlotsPerBlock];
    __shared__ uint32_t n_list;
    const uint32_t cnt = compact_block(ws.flags, total, kFilter, kFilter, list, &n_list);
    uint32_t answered = 0, handed = 0;
    WfTally tally;
    for (uint32_t i = threadIdx.x; i < cnt; i += kWfBlock) {
        const uint32_t slot = list[i];
        const uint32_t n = ws.n;
        const V3 o{ws.ray[slot], ws.ray[n + slot], ws.ray[2 * n + slot]};
        const V3 d{ws.ray[3 * n + slot], ws.ray[4 * n + slot], ws.ray[5 * n + slot]};
        const double max_t = ws.ray[6 * n + slot];
        const bool clear = scene_clear<COUNT>(search_scene(sc), o, d, max_t, tally.tc);
        const uint32_t keep = ws.flags[slot] & ~kWorkMask;
        if (clear) {
            ws.hit_t[slot] = max_t; ws.hit_id[slot] = kNone; ws.hit_rec[slot] = 0u;
            ws.flags[slot] = keep | kShade;
        } else {
            ws.flags[slot] = keep | kWalkAny;
        }
        answered += clear ? 1u : 0u;
        handed += clear ? 0u : 1u;
    }
    flush_search_stats(fstat, 0, answered, handed);
    flush_tally<COUNT>(tally, counters);
}

// ---------------------------------------------------------------- shading: RayState::advance between two traversals
template <bool COUNT>
__global__ void __launch_bounds__(kWfBlock)
shade_kernel(const __grid_constant__ DevScene sc, const __grid_constant__ DevParams P, const __grid_constant__ WfState ws,
             uint32_t* __restrict__ fb, int32_t* __restrict__ hit_ids, uint32_t* __restrict__ acc, const int tiles_x,
             const uint32_t total, const int round, DevCounters* __restrict__ counters) {
    __shared__ uint32_t list[kSlotsPerBlock];
    __shared__ uint32_t n_list;
    const uint32_t cnt = compact_block(ws.flags, total, kShade, kShade, list, &n_list);
    const bool filter_on = search_on(P, 2) && sc.global_root != 0u;
    const uint32_t n = ws.n;
    bool emitted = false;
    WfTally tally;
    for (uint32_t i = threadIdx.x; i < cnt; i += kWfBlock) {
        const uint32_t slot = list[i];
        const SlotPixel sp = slot_pixel(P, slot, tiles_x);
        const uint32_t fl = ws.flags[slot];
        RayState st;
        if (fl & kFresh) {
            st.begin(V3{P.ox, P.oy, P.oz}, sp.d);
        } else {
            st.so = V3{ws.ray[slot], ws.ray[n + slot], ws.ray[2 * n + slot]};
            st.sd = V3{ws.ray[3 * n + slot], ws.ray[4 * n + slot], ws.ray[5 * n + slot]};
            st.smax = ws.ray[6 * n + slot];
            st.hit0 = ws.hit0[slot];
            st.depth = (int)(fl & 255u); st.li = (int)((fl >> 8) & 255u); st.shadow = ((fl >> 16) & 1u) != 0u;
            st.res = 0x00FFFFFFu; st.texel = 0u; st.mid = 0;
            if (st.shadow) {
#pragma unroll
                for (int c = 0; c < 12; ++c) st.parked[c] = ws.parked[(size_t)c * n + slot];
                st.texel = ws.texel[slot]; st.mid = ws.mid[slot];
            }
            for (int l = 0; l < st.depth; ++l) {
                st.lx[l] = ws.mirror[(size_t)(4 * l) * n + slot]; st.ly[l] = ws.mirror[(size_t)(4 * l + 1) * n + slot];
                st.lz[l] = ws.mirror[(size_t)(4 * l + 2) * n + slot]; st.lk[l] = ws.mirror[(size_t)(4 * l + 3) * n + slot];
            }
        }
        Hit h;
        h.t = ws.hit_t[slot]; h.id = ws.hit_id[slot]; h.rec = ws.hit_rec[slot];
        const int depth_before = st.depth;
        if (st.advance<COUNT>(sc, P, h)) {
            if (hit_ids) hit_ids[sp.pix * 4 + sp.k] = st.hit0;
            finish_sample(acc, fb, sp, st.res);
            ws.flags[slot] = kDone;
        } else {
            ws.ray[slot] = st.so.x; ws.ray[n + slot] = st.so.y; ws.ray[2 * n + slot] = st.so.z;
            ws.ray[3 * n + slot] = st.sd.x; ws.ray[4 * n + slot] = st.sd.y; ws.ray[5 * n + slot] = st.sd.z;
            ws.ray[6 * n + slot] = st.smax;
            ws.hit0[slot] = st.hit0;
            if (st.shadow) {   // what the shadow query interrupts (a mirror ray starts a fresh hit set-up instead)
#pragma unroll
                for (int c = 0; c < 12; ++c) ws.parked[(size_t)c * n + slot] = st.parked[c];
                ws.texel[slot] = st.texel; ws.mid[slot] = st.mid;
            }
            if (st.depth > depth_before) {   // a mirror level was opened (raytracer.rs:76-88)
                const int l = st.depth - 1;
                ws.mirror[(size_t)(4 * l) * n + slot] = st.lx[l]; ws.mirror[(size_t)(4 * l + 1) * n + slot] = st.ly[l];
                ws.mirror[(size_t)(4 * l + 2) * n + slot] = st.lz[l]; ws.mirror[(size_t)(4 * l + 3) * n + slot] = st.lk[l];
            }
            const uint32_t work = st.shadow ? (filter_on ? kFilter : kWalkAny) : kWalk;
            ws.flags[slot] = (uint32_t)st.depth | ((uint32_t)st.li << 8) | (st.shadow ? 1u << 16 : 0u) | work;
            emitted = true;
        }
        if (COUNT) { tally.shadow += st.rc.shadow; tally.reflect += st.rc.reflect; }
    }
    if (emitted) ws.pending[round & 63] = 1u;
    flush_tally<COUNT>(tally, counters);
}

__global__ void wf_switch_kernel(uint32_t* __restrict__ fs, const uint32_t probe_every) {
    const uint32_t a = fs[0], h = fs[1];
    if (a + h >= 64u) fs[2] = (2ull * a >= (unsigned long long)a + h) ? 1u : 0u;
    const uint32_t sa = fs[4], sh = fs[5];
    if (sa + sh >= 64u) fs[6] = (2ull * sa >= (unsigned long long)sa + sh) ? 1u : 0u;
    fs[8] = a; fs[9] = h; fs[10] = sa; fs[11] = sh;   // what the decision was made from (rrt_scene_search_state)
    fs[0] = fs[1] = fs[4] = fs[5] = 0u;
    const uint32_t frames = ++fs[3];
    if (frames % probe_every == 0u) { fs[2] = 1u; fs[6] = 1u; }
}

}  // namespace

// ---------------------------------------------------------------- host side
void WaveBuffers::release() {
    if (base) cudaFree(base);
    base = nullptr; cap_slots = 0; bytes = 0;
}
WaveBuffers::~WaveBuffers() {
    release();
    if (h_cnt) cudaFreeHost(h_cnt);
    h_cnt = nullptr;
}

static cudaError_t wave_ensure(WaveBuffers* wb, uint32_t n_slots, int mirror_levels) {
    if (wb->cap_slots >= n_slots && wb->mirror_levels >= mirror_levels && wb->base) return cudaSuccess;
    const int levels = mirror_levels > wb->mirror_levels ? mirror_levels : wb->mirror_levels;
    wb->release();
    // per slot: ray 56 + hit 16 + parked 96 + 5 words 20 + 32 per mirror level
    const size_t n = ((size_t)n_slots + 63) & ~(size_t)63;
    const size_t per = 56 + 16 + 96 + 20 + (size_t)32 * levels;
    const size_t bytes = n * per + 256;
    cudaError_t e = cudaMalloc(&wb->base, bytes);
    if (e != cudaSuccess) return e;
    if (!wb->h_cnt && (e = cudaMallocHost((void**)&wb->h_cnt, 64)) != cudaSuccess) return e;
    wb->cap_slots = (uint32_t)n;
    wb->mirror_levels = levels;
    wb->bytes = bytes;
    return cudaSuccess;
}

static WfState wave_state(const WaveBuffers& wb) {
    WfState ws{};
    const size_t n = wb.cap_slots;
    char* p = static_cast<char*>(wb.base);
    auto take = [&p](size_t bytes) { char* q = p; p += bytes; return q; };
    ws.n = (uint32_t)n;
    ws.pending = reinterpret_cast<uint32_t*>(take(256));
    ws.ray = reinterpret_cast<double*>(take(n * 56));
    ws.hit_t = reinterpret_cast<double*>(take(n * 8));
    ws.parked = reinterpret_cast<double*>(take(n * 96));
    ws.mirror = reinterpret_cast<double*>(take(n * 32 * (size_t)wb.mirror_levels));
    ws.hit_id = reinterpret_cast<uint32_t*>(take(n * 4));
    ws.hit_rec = reinterpret_cast<uint32_t*>(take(n * 4));
    ws.texel = reinterpret_cast<uint32_t*>(take(n * 4));
    ws.res = reinterpret_cast<uint32_t*>(take(n * 4));
    ws.hit0 = reinterpret_cast<int32_t*>(take(n * 4));
    ws.mid = reinterpret_cast<int32_t*>(take(n * 4));
    ws.flags = reinterpret_cast<uint32_t*>(take(n * 4));
    return ws;
}

template <bool COUNT>
static cudaError_t wavefront_frame(const DevScene& sc, const DevParams& P, uint32_t* d_fb, int32_t* d_hit, uint32_t* d_acc,
                                   DevCounters* d_counters, cudaStream_t stream, WaveBuffers* wb) {
    const int tiles_x = (P.W + kTileW - 1) / kTileW, tiles_y = (P.n_rows_local + kTileH - 1) / kTileH;
    const unsigned long long total64 = (unsigned long long)tiles_x * tiles_y * (kTileW * kTileH * 4);
    if (total64 == 0) return cudaSuccess;
    if (total64 > 0xFFFFFFFFull - 2 * kSlotsPerBlock) return cudaErrorInvalidValue;   // > 2^32 samples: render it in shards
    const uint32_t total = (uint32_t)total64;
    cudaError_t e;
    if ((e = wave_ensure(wb, total, wb->scene_has_mirrors ? P.max_depth : 0)) != cudaSuccess) return e;
    const WfState ws = wave_state(*wb);
    const size_t npix = (size_t)P.n_rows_local * P.W;
    if ((e = cudaMemsetAsync(d_acc, 0, npix * sizeof(uint32_t), stream)) != cudaSuccess) return e;
    if ((e = cudaMemsetAsync(ws.pending, 0, 256, stream)) != cudaSuccess) return e;
    uint32_t* fstat = (P.search_mode == 1 && !COUNT) ? P.fstat : nullptr;
    const unsigned grid_cam = (unsigned)((total + kWfBlock - 1) / kWfBlock);
    const unsigned grid_q = (unsigned)((total + kSlotsPerBlock - 1) / kSlotsPerBlock);
    camera_kernel<COUNT><<<grid_cam, kWfBlock, 0, stream>>>(sc, P, ws, d_fb, d_hit, d_acc, tiles_x, total, fstat, d_counters);
    walk_kernel<COUNT><<<grid_q, kWfBlock, 0, stream>>>(sc, P, ws, tiles_x, total, d_counters);   // camera rays handed over
    int n_point = 0;
    for (int i = 0; i < P.n_lights; ++i) n_point += P.lights[i].kind == RRT_LIGHT_POINT ? 1 : 0;
    const int per_level = n_point + 1;                        // shade rounds one level of rays needs
    const int max_rounds = per_level * (P.max_depth + 1);
    // A scene without mirrors has one level of rays: per_level rounds finish every sample, and the call stays
    // asynchronous.  With mirrors the number of levels a frame needs is only known on the device: after every level the
    // host reads whether the level's last shading sent more rays out (4 bytes, a stream synchronisation).  A live sample
    // sends a ray out in EVERY shade round until it is finished (each ray is traced within its round), so "the last
    // round sent nothing" means no sample is left, whatever level each one had reached.
    int round = 0;
    while (round < max_rounds) {
        for (int k = 0; k < per_level && round < max_rounds; ++k, ++round) {
            shade_kernel<COUNT><<<grid_q, kWfBlock, 0, stream>>>(sc, P, ws, d_fb, d_hit, d_acc, tiles_x, total, round, d_counters);
            filter_kernel<COUNT><<<grid_q, kWfBlock, 0, stream>>>(sc, ws, total, fstat, d_counters);
            walk_kernel<COUNT><<<grid_q, kWfBlock, 0, stream>>>(sc, P, ws, tiles_x, total, d_counters);
        }
        if (wb->mirror_levels == 0) break;
        if ((e = cudaMemcpyAsync(wb->h_cnt, ws.pending + ((round - 1) & 63), sizeof(uint32_t), cudaMemcpyDeviceToHost, stream)) != cudaSuccess) return e;
        if ((e = cudaStreamSynchronize(stream)) != cudaSuccess) return e;
        if (wb->h_cnt[0] == 0u) break;
    }
    if (fstat) wf_switch_kernel<<<1, 1, 0, stream>>>(P.fstat, 32u);
    return cudaGetLastError();
}

cudaError_t launch_wavefront(const DevScene& sc, const DevParams& P, uint32_t* d_fb, int32_t* d_hit, uint32_t* d_acc,
                             DevCounters* d_counters, cudaStream_t stream, WaveBuffers* wb) {
    return d_counters ? wavefront_frame<true>(sc, P, d_fb, d_hit, d_acc, d_counters, stream, wb)
                      : wavefront_frame<false>(sc, P, d_fb, d_hit, d_acc, nullptr, stream, wb);
}

}  // namespace rrt
