// hard_fwd.cuh -- deterministic hard voxelization of a batch of sweeps (+ optional perturbation and
// HardSimpleVFE mean) as ONE persistent cooperative kernel with ONE grid-wide barrier, sm_100a.
//
// Replaces mmcv `hard_voxelize_forward` with a parallel formulation that gives the SAME
// result as the serial CPU walk.  At the named sizes the path moves ~30-200 MB, i.e. 5-30 us at HBM
// speed, and the round-1 kernel (six phases, five grid barriers, ~8 scattered L2 transactions per point)
// spent half of its time waiting at barriers and the rest in chains of dependent L2 round trips.  This
// version has two phases:
//
//   A insert   per point, in point order inside 256-point tiles dealt round-robin to the CTAs:
//              [perturb] -> cell id -> ONE blind 16-byte compare-and-swap of {0, 0} -> {(cell + 1) << 32 | point
//              index, 0} on the cell's table entry {key, cnt, slot0}: it either claims the entry (the point is the
//              first of its cell so far) or returns the whole entry.  The key only ever decreases (atomicMin), so it
//              ends up holding the FIRST point of the cell.  Whoever is not first -- the arriving point or the one
//              it displaces -- "arrives": one atomicAdd on the entry's counter gives it one of the cell's P-1
//              arrival slots (slot 0 in the entry, the others in the entry's row of `lists`), unordered; only a
//              cell with more than P points overflows into a list of P-1 levels kept by atomicMax (level k ends
//              up with the k-th earliest overflow point regardless of interleaving; late points are turned away by
//              one load of the last level).  A point that becomes "first so far" adds 1 to its warp's byte of the
//              tile's 8-byte counter word and to the counter of the tile's group of 32 tiles; a displaced point gets
//              its 1 taken back.  After the phase the counters hold the exact number of first points per warp /
//              tile / group without a separate flag pass.
//   -- grid barrier --
//   B emit     every CTA scans the (few) group counters in shared memory; per tile, one 8-byte load per lane
//              gives the first points of the tile's earlier group members, so a first point's voxel index
//              (= mmcv's order of first occurrence) is known after one round trip: group prefix + earlier
//              tiles + earlier warps + earlier lanes.  Candidates re-read their entry (key.lo == index: first),
//              the first point of a cell zeroes the entry again, and -- when its voxel index is below
//              max_voxels -- emits the voxel itself: its own row (a coalesced re-read), then the earliest P-1 of
//              the cell's arrivals and overflow levels in point order (a selection over <= 2(P-1) words); zero
//              padding, coors, num_points, point2voxel/slot and the HardSimpleVFE mean (sum in slot order, then
//              one division) in the same pass.
//
// The table, the rows, the overflow lists and the counters are all-zero between calls ("clean"): phase B restores
// every word phase A touched (each entry, row and list is zeroed by the cell's first point, the counters by the
// last CTA to finish), so no call pays for clearing.  A 64-bit magic + extent in the workspace header tells a
// clean workspace from a fresh or foreign-written one; the latter takes a slow path (zero-fill + one more grid
// barrier) once.  Scattered L2 transactions per in-range point: 1 atomic (insert), 1 load + 1 store (emit), plus
// one atomicAdd + one store for points that are not first -- ~3 instead of ~8.
#pragma once
#include <cooperative_groups.h>

#include "common.cuh"

namespace cg = cooperative_groups;

namespace rpc {

constexpr int kPT = 768;                        // threads per CTA of the persistent kernel
constexpr int kTile = 256;                      // points per tile: the unit that is dealt to the CTAs
constexpr int kSub = kPT / kTile;               // tiles a CTA works on per iteration (8 warps each)
constexpr int kCtasPerSM = 2;                   // resident CTAs per SM (register cap 32 at 1024 threads)
constexpr int kGrp = 32;                        // tiles per counter group (8192 points)
constexpr int kMaxGroups = 8192;                // group counters are scanned in shared memory (32 KB)
constexpr int kL1Stride = 64;                   // ints between two group counters: 256 bytes, so that the REDs of one
                                                // iteration (256 per counter) spread over L2 slices instead of one line
constexpr int kTr = 32;                         // tiles per team and phase whose (sweep, range) is looked up once, at kernel start
constexpr int kQCap = 1024;                     // queue of in-range points per team: a round being read (256) + < 256 left + one tile pushed
constexpr int kFastD = 15;                      // P-1 up to which a voxel's arrivals are sorted in shared memory
constexpr uint32_t kWBase = 0x40000000u;        // level value of point i: kWBase - i  (> 0, larger = earlier)
constexpr unsigned long long kWsMagic = 0x325F4E4C43325052ull;   // "RP2CLN_2"
constexpr size_t kWsHeaderBytes = 256;
constexpr size_t kWsStampBytes = 8 * 16 * 4096;     // per-CTA stamps + per-tile records of the RPC_DBG=64 experiment

struct __align__(16) Entry {
    unsigned long long key;                     // 0 = empty; ((cell + 1) << 32) | index of the first point so far
    uint32_t cnt;                               // points of the cell that arrived after its (current) first point
    uint32_t slot0;                             // W of arrival 0 (arrivals 1..P-2 sit in the entry's row of `lists`)
};

struct WsHeader {
    unsigned long long magic;                   // kWsMagic when the clean region is all-zero up to clean_hi
    unsigned long long clean_hi;                // byte offset: [clean_lo, clean_hi) is zero between calls
    unsigned long long ws_bytes;                // size the header was written for (the dirty tail sits at the end)
    unsigned int done;                          // CTAs that finished phase B (reset by the last one)
    unsigned int pad;                           // (probe counter of the RPC_DBG=16 experiment)
    unsigned int next_a, next_b;                // tile counters of the two phases (reset by the last CTA)
};

struct FwdParams {
    const float *pts, *raw;                     // [ntotal, C]
    WsHeader *hdr;
    Entry *tab;                                 // [tab_total]
    uint32_t *hslot;                            // [ntotal]   1 + table entry of a candidate first point, else 0
    unsigned long long *l0;                     // [ntiles]   first points per warp of the tile, one byte each
    int *l1;                                    // [ng1 * kL1Stride]  first points per group of kGrp tiles
    uint32_t *lists;                            // [tab_total * LS]  per entry: W of arrivals 1..P-2, then the id of its overflow list
    uint32_t *ovf;                              // [ntotal * LS]     overflow lists (cells with more than P points): P-1 levels
    float *voxels; int *coors; int *num_points; int *voxel_num; int *p2v; int *p2s; float *feats;
    unsigned long long *stamps;                 // [gridDim][16] %globaltimer per CTA and phase, or NULL
    long long tab_total;
    unsigned long long clean_lo, front_end, back_off, ws_bytes;
    int ntiles, ng1, LS;
    int DS;                                     // words of shared-memory sort scratch per thread in phase B (odd, >= P-1), 0: none
    int scan_words;                             // ints of shared memory in front of the scratch (group prefix)
    int pf_mode;                                // table prefetch: 0 none, 1 per line, 2 bulk (TMA)
    int dbg;                                    // RPC_DBG bits (timing experiments; 1-4 give wrong results): 1 no arrivals, 2 no counters, 4 no map init, 8 no point prefetch, 16 count probes, 64 per-tile records, 256 plain (not evict_last) entry zeroing
    GridDesc gd;
    BatchDesc bd;
    PerturbDesc pd;
};

// read x,y,z of row i (and the perturbation when PERTURB) -- C == 4 uses one 16-byte load
template <int C, bool PERTURB>
__device__ __forceinline__ void load_xyz(const float *__restrict__ pts, const float *__restrict__ raw,
                                         const PerturbDesc &pd, int64_t i, float &x, float &y, float &z)
{
    if constexpr (C == 4) {
        float4 v = __ldg(reinterpret_cast<const float4 *>(pts) + i);
        x = v.x; y = v.y; z = v.z;
        if constexpr (PERTURB) {
            float4 r = __ldg(reinterpret_cast<const float4 *>(raw) + i);
            x = __fadd_rn(x, perturb_delta(r.x, pd.eb[0], pd.fb[0]));
            y = __fadd_rn(y, perturb_delta(r.y, pd.eb[1], pd.fb[1]));
            z = __fadd_rn(z, perturb_delta(r.z, pd.eb[2], pd.fb[2]));
        }
    } else {
        const float *p = pts + i * C;
        x = __ldg(p); y = __ldg(p + 1); z = __ldg(p + 2);
        if constexpr (PERTURB) {
            const float *r = raw + i * C;
            x = __fadd_rn(x, perturb_delta(__ldg(r), pd.eb[0], pd.fb[0]));
            y = __fadd_rn(y, perturb_delta(__ldg(r + 1), pd.eb[1], pd.fb[1]));
            z = __fadd_rn(z, perturb_delta(__ldg(r + 2), pd.eb[2], pd.fb[2]));
        }
    }
}

// the whole (perturbed) row of point i
template <int C, bool PERTURB>
__device__ __forceinline__ void load_row(const FwdParams &p, int64_t i, float (&x)[C])
{
    if constexpr (C == 4) {
        const float4 t4 = __ldg(reinterpret_cast<const float4 *>(p.pts) + i);
        x[0] = t4.x; x[1] = t4.y; x[2] = t4.z; x[3] = t4.w;
        if constexpr (PERTURB) {
            const float4 d = __ldg(reinterpret_cast<const float4 *>(p.raw) + i);
            const float dv[4] = {d.x, d.y, d.z, d.w};
#pragma unroll
            for (int c = 0; c < 4; ++c) x[c] = __fadd_rn(x[c], perturb_delta(dv[c], p.pd.eb[c], p.pd.fb[c]));
        }
    } else {
#pragma unroll
        for (int c = 0; c < C; ++c) x[c] = __ldg(p.pts + i * C + c);
        if constexpr (PERTURB) {
#pragma unroll
            for (int c = 0; c < C; ++c)
                x[c] = __fadd_rn(x[c], perturb_delta(__ldg(p.raw + i * C + c), p.pd.eb[c], p.pd.fb[c]));
        }
    }
}

__device__ __forceinline__ void stamp(const FwdParams &p, int k)
{
    if (p.stamps && threadIdx.x == 0) {
        unsigned long long t;
        asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t));
        p.stamps[blockIdx.x * 16 + k] = t;
    }
}

struct TileRange { int b, begin, end; };
__device__ __forceinline__ TileRange tile_range(const FwdParams &p, int tile)
{
    TileRange r;
    r.b = tile_to_sweep(p.bd, tile);
    r.begin = p.bd.pt_begin[r.b] + (tile - p.bd.tile_begin[r.b]) * kTile;
    r.end = min(r.begin + kTile, p.bd.pt_begin[r.b + 1]);
    return r;
}

// first table entry probed for a cell: multiplicative hash, then a multiply-shift range reduction
// (table sizes are not powers of two: 1.5 entries per point)
__device__ __forceinline__ uint32_t probe_start(uint32_t cell, uint32_t size)
{
    uint32_t h = cell * 0x9E3779B1u;
    h ^= h >> 15;
    h *= 0x85EBCA77u;
    return __umulhi(h, size);
}

// sum of the 8 per-warp byte counts of a tile (each <= 32)
__device__ __forceinline__ int tile_first_points(unsigned long long w8)
{
    return (int)(__vsadu4((uint32_t)w8, 0u) + __vsadu4((uint32_t)(w8 >> 32), 0u));
}

// ---- TMA (bulk async copy) + mbarrier helpers for the point-tile pipeline of phase A (C == 4: a warp's rows are
// a contiguous 16-byte-aligned range, so one cp.async.bulk per array brings them into shared memory while the
// previous iteration is still probing the table)
__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t count)
{
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint32_t bar)
{
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bar) : "memory");
}
__device__ __forceinline__ void mbar_arrive_expect_tx(uint32_t bar, uint32_t bytes)
{
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
}
__device__ __forceinline__ void tma_load_1d(uint32_t dst, const void *src, uint32_t bytes, uint32_t bar)
{
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                 ::"r"(dst), "l"(src), "r"(bytes), "r"(bar) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity)
{
    uint32_t ok;
    do {
        asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}"
                     : "=r"(ok) : "r"(bar), "r"(parity) : "memory");
    } while (!ok);
}

// blind claim of an entry: 16-byte compare-and-swap of {0, 0} -> {key, 0}; returns what was there -- for an entry
// that already belongs to a cell that is its key AND its counter / slot-0 words in the same round trip
// (SASS: ATOMG.E.CAS.128; measured 81 G/s against 100 G/s for the 64-bit CAS, tools/micro/atomics2.cu)
__device__ __forceinline__ uint4 cas_entry(Entry *e, unsigned long long key)
{
    unsigned long long o0, o1;
    asm volatile("{\n\t.reg .b128 c, v, o;\n\tmov.b128 c, {%3, %3};\n\tmov.b128 v, {%4, %3};\n\t"
                 "atom.global.relaxed.gpu.cas.b128 o, [%2], c, v;\n\tmov.b128 {%0, %1}, o;\n\t}"
                 : "=l"(o0), "=l"(o1) : "l"(e), "l"(0ull), "l"(key) : "memory");
    return make_uint4((uint32_t)o0, (uint32_t)(o0 >> 32), (uint32_t)o1, (uint32_t)(o1 >> 32));
}

// TMA bulk prefetch of [p, p + bytes) into L2 (16-byte aligned, bytes % 16 == 0): one instruction per chunk
__device__ __forceinline__ void bulk_prefetch_l2(const void *ptr, uint32_t bytes)
{
    asm volatile("cp.async.bulk.prefetch.L2.global [%0], %1;" ::"l"(ptr), "r"(bytes) : "memory");
}

// The table lives in L2 ACROSS calls: the store that zeroes an entry at the end of a call carries an L2::evict_last
// cache policy, so whatever runs between two calls (bench.py overwrites 512 MiB) evicts other lines first and the
// next call's atomics hit L2 instead of fetching a 32-byte sector from DRAM each (tools/micro/atomics4.cu: 313 k
// CASes 9.2 us warm, 14.4 us cold, 10.2 us after an overwrite when the zeroing stores carried the policy; ptxas
// rejects the hint on atom itself).  17 MB (KITTI batch) / 29 MB (nuScenes) of the 126 MB L2.
__device__ __forceinline__ void st_zero16_keep(void *ptr)
{
    unsigned long long pol;
    asm volatile("createpolicy.fractional.L2::evict_last.b64 %0, 1.0;" : "=l"(pol));
    asm volatile("st.global.L2::cache_hint.v4.u32 [%0], {%1,%1,%1,%1}, %2;" ::"l"(ptr), "r"(0u), "l"(pol) : "memory");
}
__device__ __forceinline__ void st_zero4_keep(void *ptr)
{
    unsigned long long pol;
    asm volatile("createpolicy.fractional.L2::evict_last.b64 %0, 1.0;" : "=l"(pol));
    asm volatile("st.global.L2::cache_hint.u32 [%0], %1, %2;" ::"l"(ptr), "r"(0u), "l"(pol) : "memory");
}

__device__ __forceinline__ uint4 ld_entry(const Entry *e)
{
    uint4 v;
    asm volatile("ld.global.cg.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(e) : "memory");
    return v;
}

// zero `n` floats at `out` (4-byte aligned) with the lanes of one warp: scalar head up to the first 16-byte
// boundary, 16-byte body, scalar tail
__device__ __forceinline__ void warp_zero_fill(float *out, int n, int lane)
{
    int head = (int)(((16u - (uint32_t)(reinterpret_cast<uintptr_t>(out) & 15u)) & 15u) >> 2);
    head = head < n ? head : n;
    if (lane < head) out[lane] = 0.0f;
    float4 *b4 = reinterpret_cast<float4 *>(out + head);
    const int n4 = (n - head) >> 2;
    const float4 z4 = make_float4(0.f, 0.f, 0.f, 0.f);
    for (int q = lane; q < n4; q += 32) b4[q] = z4;
    for (int q = head + (n4 << 2) + lane; q < n; q += 32) out[q] = 0.0f;
}

// (b, first point, end, tile) of a tile, or tile = -1 past the end.  Called by all 32 lanes of a warp: lane b tests
// one sweep boundary, so the sweep of a tile costs one parameter load per lane and a ballot (a serial search by one
// thread per team was ~1 us per tile and phase).
__device__ __forceinline__ int4 warp_range(const FwdParams &p, int tile, int lane)
{
    const bool ge = lane + 1 < p.bd.B && tile >= p.bd.tile_begin[lane + 1 < RPC_MAX_BATCH ? lane + 1 : RPC_MAX_BATCH];
    const int b = __popc(__ballot_sync(0xFFFFFFFFu, ge));
    if (tile < 0 || tile >= p.ntiles) return make_int4(0, 0, 0, -1);
    const int begin = p.bd.pt_begin[b] + (tile - p.bd.tile_begin[b]) * kTile;
    return make_int4(b, begin, min(begin + kTile, p.bd.pt_begin[b + 1]), tile);
}

// Order in which phase B takes the tiles: round-robin over the sweeps (k-th tile of sweep 0, of sweep 1, ...), so that
// the tiles at the START of every sweep -- where the cells are first seen, i.e. where the voxels are emitted -- come
// first and spread evenly over the teams: longest jobs first.  j = 0, 1, 2, ... -> tile (a bijection; ragged sweeps are
// skipped once they run out).  Warp-parallel: lane b holds the tile count of sweep b.
__device__ __forceinline__ int emit_order(const FwdParams &p, int j, int lane)
{
    const int B = p.bd.B;
    const int nt = lane < B ? p.bd.tile_begin[lane + 1] - p.bd.tile_begin[lane] : 0x7FFFFFFF;
    int tmin = nt;
#pragma unroll
    for (int d = 16; d > 0; d >>= 1) tmin = min(tmin, __shfl_xor_sync(0xFFFFFFFFu, tmin, d));
    if (j < 0 || j >= p.ntiles) return -1;
    if (j < tmin * B) return p.bd.tile_begin[j % B] + j / B;
    int left = j - tmin * B;
    for (int k = tmin;; ++k) {                                       // (ragged tail: layer k holds one tile of every sweep longer than k)
        const uint32_t alive = __ballot_sync(0xFFFFFFFFu, lane < B && nt > k);
        const int na = __popc(alive);
        if (left < na) return p.bd.tile_begin[__fns(alive, 0, left + 1)] + k;
        left -= na;
        if (na == 0) return -1;
    }
}

// serial versions (one thread), used once per tile at kernel start to fill the per-team tile tables
__device__ __forceinline__ int4 serial_range(const FwdParams &p, int tile)
{
    if (tile < 0 || tile >= p.ntiles) return make_int4(0, 0, 0, -1);
    const TileRange tr = tile_range(p, tile);
    return make_int4(tr.b, tr.begin, tr.end, tile);
}
__device__ __forceinline__ int serial_emit_order(const FwdParams &p, int j)
{
    if (j < 0 || j >= p.ntiles) return -1;
    const int B = p.bd.B;
    int tmin = 0x7FFFFFFF;
    for (int b = 0; b < B; ++b) tmin = min(tmin, p.bd.tile_begin[b + 1] - p.bd.tile_begin[b]);
    if (j < tmin * B) return p.bd.tile_begin[j % B] + j / B;
    int left = j - tmin * B;
    for (int k = tmin;; ++k) {
        int alive = 0;
        for (int b = 0; b < B; ++b) alive += (p.bd.tile_begin[b + 1] - p.bd.tile_begin[b]) > k;
        if (left < alive) {
            for (int b = 0; b < B; ++b)
                if ((p.bd.tile_begin[b + 1] - p.bd.tile_begin[b]) > k && left-- == 0) return p.bd.tile_begin[b] + k;
        }
        left -= alive;
        if (alive == 0) return -1;
    }
}

// barrier of one team (the 8 warps = 256 threads that work on one tile); ids 1..kSub, 0 is __syncthreads
__device__ __forceinline__ void team_sync(int team)
{
    switch (team) {                                                  // immediate ids: only five barriers are reserved
        case 0: asm volatile("bar.sync 1, 256;" ::: "memory"); break;
        case 1: asm volatile("bar.sync 2, 256;" ::: "memory"); break;
        case 2: asm volatile("bar.sync 3, 256;" ::: "memory"); break;
        default: asm volatile("bar.sync 4, 256;" ::: "memory"); break;
    }
}

template <int C, bool PERTURB>
__global__ void __launch_bounds__(kPT, kCtasPerSM)
k_hard_fwd(const __grid_constant__ FwdParams p)
{
    cg::grid_group grid = cg::this_grid();
    extern __shared__ __align__(16) unsigned char dyn_sm[];     // A: TMA stages / B: group prefix
    __shared__ int sm[33];
    __shared__ int s_voff[RPC_MAX_BATCH + 1];                   // first voxel row of each sweep, then M
    __shared__ int s_base[RPC_MAX_BATCH];                       // first points in the groups before the sweep's first
    __shared__ __align__(8) unsigned long long s_bar[2 * (kPT / 32)];   // A (C == 4): one mbarrier per warp and stage
    __shared__ int s_qcell[kSub][kQCap], s_qidx[kSub][kQCap];   // per team: queue of in-range points (cell, index, sweep)
    __shared__ unsigned char s_qsb[kSub][kQCap];
    __shared__ uint32_t s_qpush[kSub][3];                       // points pushed in iteration it (slot it % 3; reset two iterations later)
    __shared__ int s_flag;
    __shared__ int4 s_tr[2][kSub][kTr];                         // [phase][team][k]: the team's k-th tile (sweep, first, end, tile)
    const int tid = threadIdx.x;
    const int P = p.gd.P, V = p.gd.V, LS = p.LS;
    const int nct = gridDim.x;
    const int wlane = tid & 31, wid = tid >> 5;
    const int team = tid >> 8, tl = tid & (kTile - 1);           // a team = 8 warps = one tile at a time

    stamp(p, 0);
    if (p.stamps && threadIdx.x == 0) { p.stamps[blockIdx.x * 16 + 8] = 0ull; p.stamps[blockIdx.x * 16 + 9] = 0ull; }
    if (p.stamps && threadIdx.x == 0) {
        unsigned int smid;
        asm volatile("mov.u32 %0, %smid;" : "=r"(smid));
        p.stamps[blockIdx.x * 16 + 7] = smid;
    }
    // is the workspace clean (all-zero table, rows, lists and counters, left so by the previous call)?
    if (tid == 0) {
        const unsigned long long *h = reinterpret_cast<const unsigned long long *>(p.hdr);
        const unsigned long long magic = __ldcg(h), chi = __ldcg(h + 1), wsb = __ldcg(h + 2);
        s_flag = (magic == kWsMagic && wsb == p.ws_bytes && chi >= p.front_end) ? 1 : 0;
    }
    // Tiles are handed out dynamically: a team's first tile is fixed (team t of CTA c starts with tile c + t * G),
    // every further one comes from a global counter that the team's leader increments one iteration ahead.  Tiles differ by 6x in cost (an in-range tile waits for its atomics, an
    // out-of-range one does not), so a static deal left most CTAs waiting at the barrier for the slowest.
    const int first = (int)blockIdx.x + team * nct, stride = kSub * nct;   // this team's tiles: first + k * stride
    // the (sweep, range) of every tile this CTA will touch, looked up once: doing it per iteration in every warp
    // (index arithmetic, searches over the sweep boundaries in parameter space) cost 3 us per iteration and phase
    if (tid < 2 * kSub * kTr) {
        const int ph = tid / (kSub * kTr), t = (tid / kTr) % kSub, k = tid % kTr;
        const int j = (int)blockIdx.x + t * nct + k * stride;
        s_tr[ph][t][k] = serial_range(p, ph ? serial_emit_order(p, j) : j);
    }
    __syncthreads();
    auto tile_a = [&](int k) -> int4 { return k < kTr ? s_tr[0][team][k] : warp_range(p, first + k * stride, wlane); };
    auto tile_b = [&](int k) -> int4 { return k < kTr ? s_tr[1][team][k] : warp_range(p, emit_order(p, first + k * stride, wlane), wlane); };
    int4 cur = tile_a(0);
    const bool was_clean = s_flag != 0;
    if (p.stamps && tid == 0) p.stamps[blockIdx.x * 16 + 15] = was_clean ? 1ull : 0ull;
    // (the staging pipeline of phase A is set up here, and the rows of the first tile are requested first thing)
    constexpr bool kTma = (C == 4);
    constexpr int kArr = PERTURB ? 2 : 1;                            // arrays staged per warp
    const uint32_t bar_sa = (uint32_t)__cvta_generic_to_shared(s_bar) + 8u * wid;          // + 8 * 32 * stage
    const uint32_t stage_sa = (uint32_t)__cvta_generic_to_shared(dyn_sm) + (uint32_t)(wid * kArr * 512);   // + stage
    constexpr uint32_t kStageBytes = (kPT / 32) * kArr * 512;
    auto issue_rows = [&](const int4 &t, int st) {                   // called by lane 0 of every warp
        const uint32_t bar = bar_sa + 8u * (kPT / 32) * st;
        int r0 = 0, r1 = 0;
        if (t.w >= 0) {
            r0 = t.y + (wid & 7) * 32;
            r1 = min(r0 + 32, t.z);
        }
        if (r1 <= r0) { mbar_arrive(bar); return; }
        const uint32_t bytes = (uint32_t)(r1 - r0) * 16u;
        const uint32_t dst = stage_sa + kStageBytes * st;
        mbar_arrive_expect_tx(bar, bytes * kArr);
        tma_load_1d(dst, reinterpret_cast<const float4 *>(p.pts) + r0, bytes, bar);
        if constexpr (PERTURB) tma_load_1d(dst + 512, reinterpret_cast<const float4 *>(p.raw) + r0, bytes, bar);
    };
    if constexpr (kTma) {
        if (wlane == 0) {
            mbar_init(bar_sa, 1);
            mbar_init(bar_sa + 8u * (kPT / 32), 1);
            asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
            issue_rows(cur, 0);
        }
        __syncwarp();
    }
    // pull the team's later tiles into L2 now (static deal: tiles first + k * 4 G): their TMA loads then hit L2 instead
    // of paying the HBM latency once per iteration
    if (tl < 32 && !(p.dbg & 8)) {
        for (int t = first + stride; t < p.ntiles; t += stride) {
            const int4 r = tile_a((t - first) / stride);
            if (wlane != 0) continue;
            const size_t b0 = ((size_t)r.y * C * 4) & ~size_t(15), b1 = ((size_t)r.z * C * 4 + 15) & ~size_t(15);
            bulk_prefetch_l2(reinterpret_cast<const char *>(p.pts) + b0, (uint32_t)(b1 - b0));
            if constexpr (PERTURB) bulk_prefetch_l2(reinterpret_cast<const char *>(p.raw) + b0, (uint32_t)(b1 - b0));
        }
    }
    if (was_clean) {
        // (experiment knob: stream the table into L2 while the first rows travel; off by default, the table stays
        // L2-resident across calls through the evict_last policy of the stores that zero it)
        if (p.pf_mode == 1) {
            const char *t = reinterpret_cast<const char *>(p.tab);
            const long long nl = (p.tab_total * (long long)sizeof(Entry) + 127) >> 7;
            for (long long i = (long long)blockIdx.x * kPT + tid; i < nl; i += (long long)nct * kPT)
                asm volatile("prefetch.global.L2 [%0];" ::"l"(t + (i << 7)));
        } else if (p.pf_mode == 2 && wlane == 0) {
            constexpr long long kChunk = 16384;
            const char *t = reinterpret_cast<const char *>(p.tab);
            const long long nb = p.tab_total * (long long)sizeof(Entry);
            for (long long o = ((long long)blockIdx.x * (kPT / 32) + wid) * kChunk; o < nb; o += (long long)nct * (kPT / 32) * kChunk)
                bulk_prefetch_l2(t + o, (uint32_t)(nb - o < kChunk ? nb - o : kChunk));
        }
    } else {
        // slow path (first use of this workspace, or someone else wrote to it): zero the clean region
        uint4 *z = reinterpret_cast<uint4 *>(reinterpret_cast<char *>(p.hdr) + p.clean_lo);
        const long long n4 = (long long)((p.front_end - p.clean_lo) >> 4);
        for (long long i = (long long)blockIdx.x * kPT + tid; i < n4; i += (long long)nct * kPT)
            z[i] = make_uint4(0u, 0u, 0u, 0u);
        if (blockIdx.x == 0 && tid == 0) { p.hdr->done = 0u; p.hdr->next_a = 0u; p.hdr->next_b = 0u; }
        grid.sync();
    }
    stamp(p, 1);

    // ------------------------------------------------------------------------------- A insert
    // Two stages per team.  Stage 1, tile by tile: rows -> [perturb] -> cell; the points that are in range are pushed
    // into the team's queue in shared memory (the others are done).  Stage 2, whenever the queue holds a full round
    // of 256 points (and once more at the end): every thread takes one queued point through the table.  An
    // out-of-range tile costs ~0.8 us, a round of atomics ~4 us whatever the number of its points, so rounds are
    // only ever run full, and a team that runs rounds simply grabs fewer tiles: the SMs stay evenly loaded although
    // a LiDAR sweep's in-range points come in long runs (KITTI: 56 % of the tiles are entirely out of range).
    uint32_t qhead = 0, qtail = 0;                                   // (team-uniform) the queue holds positions [qhead, qtail) mod kQCap
    if (tl < 3) s_qpush[team][tl] = 0u;
    team_sync(team);
    for (int it = 0;; ++it) {
        const bool last = cur.w < 0;                                 // (the same for the whole team) no tile left: drain
        int4 nxt = cur;
        if (!last) {
            nxt = tile_a(it + 1);
            if constexpr (kTma) {
                // every lane of this warp is past its reads of the other stage (__syncwarp below): refill it
                if (nxt.w >= 0 && wlane == 0) issue_rows(nxt, (it + 1) & 1);
            }
            // ---- stage 1: this tile's cells
            const int idx = cur.y + tl < cur.z ? cur.y + tl : -1;
            int cell = -1;
            if constexpr (kTma) mbar_wait(bar_sa + 8u * (kPT / 32) * (it & 1), (uint32_t)(it >> 1) & 1u);
            if (idx >= 0) {
                float x, y, z;
                if constexpr (kTma) {
                    const float4 *sp = reinterpret_cast<const float4 *>(dyn_sm) +
                                       ((it & 1) * (kPT / 32) + wid) * kArr * 32 + wlane;
                    const float4 v = sp[0];
                    x = v.x; y = v.y; z = v.z;
                    if constexpr (PERTURB) {
                        const float4 r = sp[32];
                        x = __fadd_rn(x, perturb_delta(r.x, p.pd.eb[0], p.pd.fb[0]));
                        y = __fadd_rn(y, perturb_delta(r.y, p.pd.eb[1], p.pd.fb[1]));
                        z = __fadd_rn(z, perturb_delta(r.z, p.pd.eb[2], p.pd.fb[2]));
                    }
                } else {
                    load_xyz<C, PERTURB>(p.pts, p.raw, p.pd, idx, x, y, z);
                }
                cell = point_cell(x, y, z, p.gd);
                if (p.p2v) p.p2v[idx] = -1;
                if (p.p2s) p.p2s[idx] = -1;
                if (cell < 0) p.hslot[idx] = 0u;                     // (in-range points get their word in stage 2)
            }
            if constexpr (kTma) __syncwarp();                        // the stage may be refilled from here on
            const uint32_t inr = __ballot_sync(0xFFFFFFFFu, cell >= 0);
            if (inr != 0u) {
                uint32_t base = 0;
                if (wlane == 0) base = atomicAdd(&s_qpush[team][it % 3], (uint32_t)__popc(inr));
                base = __shfl_sync(0xFFFFFFFFu, base, 0);
                if (cell >= 0) {
                    const uint32_t q = (qtail + base + (uint32_t)__popc(inr & ((1u << wlane) - 1u))) & (kQCap - 1);
                    s_qcell[team][q] = cell;
                    s_qidx[team][q] = idx;
                    s_qsb[team][q] = (unsigned char)cur.x;
                }
            }
            team_sync(team);                                         // the pushes of all 8 warps are in
            qtail += s_qpush[team][it % 3];
            if (tl == 0) s_qpush[team][(it + 2) % 3] = 0u;           // (slot of iteration it - 1: read by all, next used at it + 2)
        }
        const uint32_t qn = qtail - qhead;
        // ---- stage 2: one round, when it is full or when nothing more will come
        if ((qn >= (uint32_t)kTile || (last && qn > 0u)) && !(p.dbg & 1024)) {
            const uint32_t take = qn < (uint32_t)kTile ? qn : (uint32_t)kTile;
            if (p.stamps && tl == 0) {                               // (experiments) rounds and points of this CTA, time in rounds
                atomicAdd(p.stamps + blockIdx.x * 16 + 8, 1ull);
                atomicAdd(p.stamps + blockIdx.x * 16 + 9, (unsigned long long)take);
            }
            int idx = -1, cell = -1, sb = 0;
            if ((uint32_t)tl < take) {
                const uint32_t q = (qhead + (uint32_t)tl) & (kQCap - 1);
                cell = s_qcell[team][q]; idx = s_qidx[team][q]; sb = s_qsb[team][q];
            }
            qhead += take;
            const uint32_t off = p.bd.tab_off[sb], size = p.bd.tab_size[sb];
            // claim / find the cell's entry: blind 16-byte CAS along the linear probe order (at LiDAR load factors
            // ~0.3 the home entry is free or already the cell's own for most points: one L2 round trip)
            bool cand = false;                                       // first point of the cell so far
            uint32_t h = 0;
            int disp = -1;                                           // displaced "first so far"
            if (cell >= 0) {
                const uint32_t ck = (uint32_t)cell + 1u;
                const unsigned long long mine = ((unsigned long long)ck << 32) | (uint32_t)idx;
                uint32_t pos = probe_start((uint32_t)cell, size);
                uint32_t w = 0u, cnt = 0u;                           // level value of whoever arrives; snapshot of the counter
                // (bounded: a table that someone else scribbled into must not hang the GPU; with a clean table a
                // free entry always exists -- 1.5 entries per point)
                for (uint32_t probes = 0; probes < size; ++probes) {
                    Entry *e = p.tab + off + pos;
                    if (p.dbg & 16) atomicAdd(&p.hdr->pad, 1u);
                    const uint4 o = cas_entry(e, mine);
                    if ((o.x | o.y) == 0u) { cand = true; h = off + pos; break; }
                    if (o.y == ck) {
                        h = off + pos;
                        int carry = idx;
                        cnt = o.z;
                        if (o.x > (uint32_t)idx) {
                            const unsigned long long old = atomicMin(&e->key, mine);
                            if ((uint32_t)old > (uint32_t)idx) { cand = true; disp = (int)(uint32_t)old; carry = disp; }
                        }
                        w = kWBase - (uint32_t)carry;
                        break;
                    }
                    pos = pos + 1 == size ? 0u : pos + 1;
                }
                // A point that is not the first of its cell ARRIVES: it takes the next of the cell's P-1 arrival slots
                // (one atomicAdd on the entry's counter, one plain store) -- unordered; the first point sorts them
                // when it emits the voxel.  Only a cell with more than P points overflows: arrival P-1 and later
                // compete for the P-1 levels of an overflow list (every level keeps the largest W = earliest point it
                // has seen, the loser moves on; levels only grow, so a late point whose W is below the LAST level is
                // turned away by one load -- LiDAR null returns all fall into the cell of the origin).  The voxel's
                // P-1 later points are the earliest of arrivals + overflow levels.
                if (w != 0u && P > 1 && !(p.dbg & 1)) {
                    Entry *e = p.tab + h;
                    const uint32_t D = (uint32_t)P - 1u;
                    uint32_t *row = p.lists + (size_t)h * LS;
                    bool over = cnt > D;                             // (snapshot) someone overflowed already
                    if (!over) {
                        const uint32_t r = atomicAdd(&e->cnt, 1u);
                        if (r == 0u) __stcg(&e->slot0, w);
                        else if (r < D) __stcg(row + r - 1u, w);
                        else over = true;
                    }
                    if (over) {
                        bool fresh = false;
                        uint32_t lp = __ldcg(row + D - 1u);
                        if (lp == 0u) {
                            lp = atomicCAS(row + D - 1u, 0u, (uint32_t)idx + 1u);
                            if (lp == 0u) { lp = (uint32_t)idx + 1u; fresh = true; }
                        }
                        uint32_t *L = p.ovf + (size_t)(lp - 1u) * LS;
                        if (fresh || D == 1u || __ldcg(L + D - 1u) < w) {
                            for (uint32_t q = 0; q < D; ++q) {
                                const uint32_t old = atomicMax(L + q, w);
                                if (old == 0u) break;
                                w = old < w ? old : w;
                            }
                        }
                    }
                }
                p.hslot[idx] = cand ? h + 1u : 0u;
            }
            // first points so far: +1 in the byte of the point's warp in its tile's word and in the group's counter
            // (REDs, aggregated over the lanes that share the word / the counter: queued points keep their order, so
            // a warp of the round mostly covers one or two warps of one tile)
            const uint32_t cm = __ballot_sync(0xFFFFFFFFu, cand);
            if (cand && !(p.dbg & 2)) {
                const int jl = idx - p.bd.pt_begin[sb];
                const uint32_t g0 = __match_any_sync(cm, (uint32_t)(jl >> 5) | ((uint32_t)sb << 26));
                if ((g0 & ((1u << wlane) - 1u)) == 0u)
                    atomicAdd(p.l0 + p.bd.tile_begin[sb] + jl / kTile, (unsigned long long)__popc(g0) << (8 * ((jl >> 5) & 7)));
                const uint32_t g1 = __match_any_sync(cm, (uint32_t)(jl / (kTile * kGrp)) | ((uint32_t)sb << 26));
                if ((g1 & ((1u << wlane) - 1u)) == 0u)
                    atomicAdd(p.l1 + (size_t)(p.bd.g1_begin[sb] + jl / (kTile * kGrp)) * kL1Stride, __popc(g1));
            }
            if (disp >= 0) {                                         // the displaced point is not first after all
                const int jl = disp - p.bd.pt_begin[sb];
                atomicAdd(p.l0 + p.bd.tile_begin[sb] + jl / kTile, ~(1ull << (8 * ((jl >> 5) & 7))) + 1ull);
                atomicAdd(p.l1 + (size_t)(p.bd.g1_begin[sb] + jl / (kTile * kGrp)) * kL1Stride, -1);
            }
        }
        if (last) break;
        cur = nxt;
    }
    if (p.stamps) __syncthreads();
    stamp(p, 2);
    grid.sync();
    stamp(p, 3);

    // ------------------------------------------------------------------------------- B emit
    // exclusive scan of the group counters (first points in the groups before g), every CTA for itself
    int *s_E = reinterpret_cast<int *>(dyn_sm);
    uint32_t *s_sc = reinterpret_cast<uint32_t *>(dyn_sm) + p.scan_words;       // [kPT][DS] per-thread sort scratch
    {
        const int per = (p.ng1 + kPT - 1) / kPT;
        const int j0 = tid * per, j1 = min(j0 + per, p.ng1);
        int s = 0;
        for (int j = j0; j < j1; ++j) s += __ldcg(p.l1 + (size_t)j * kL1Stride);
        int tot;
        int run = block_excl_scan<kPT>(s, sm, &tot);
        for (int j = j0; j < j1; ++j) { s_E[j] = run; run += __ldcg(p.l1 + (size_t)j * kL1Stride); }
        if (tid == kPT - 1) s_E[p.ng1] = tot;
        __syncthreads();
        if (tid == 0) {
            int acc = 0;
            for (int b = 0; b < p.bd.B; ++b) {
                const int base = s_E[p.bd.g1_begin[b]];
                const int m = min(s_E[p.bd.g1_begin[b + 1]] - base, V);
                s_base[b] = base;
                s_voff[b] = acc;
                if (blockIdx.x == 0) p.voxel_num[b] = m;
                acc += m;
            }
            s_voff[p.bd.B] = acc;
            if (blockIdx.x == 0) p.voxel_num[p.bd.B] = acc;
        }
        __syncthreads();
    }
    stamp(p, 4);
    const int F = p.gd.F;
    const int PC = P * C;
    // (no team barrier in this phase: every warp walks the team's tiles on its own, 32 consecutive points of each)
    // (the tile descriptors are re-read from shared memory every iteration rather than carried in registers: this
    // phase is the register-hungry one, and a spilled register is an L2 round trip here)
    uint32_t hs;
    {
        const int4 c0 = tile_b(0);
        hs = (c0.w >= 0 && c0.y + tl < c0.z && !(p.dbg & 32768)) ? __ldcg(p.hslot + c0.y + tl) : 0u;
    }
#pragma unroll 1
    for (int it = 0;; ++it) {
        int tile, sb, idx;
        {
            const int4 c = tile_b(it);
            if (c.w < 0) break;
            tile = c.w; sb = c.x;
            idx = c.y + tl < c.z ? c.y + tl : -1;
        }
        // the next tile's entry words travel while this one is emitted
        uint32_t hs_nxt;
        {
            const int4 nxt = tile_b(it + 1);
            hs_nxt = (nxt.w >= 0 && nxt.y + tl < nxt.z && !(p.dbg & 32768)) ? __ldcg(p.hslot + nxt.y + tl) : 0u;
        }
        unsigned long long t_tile0 = 0;
        if (p.stamps && (p.dbg & 64) && tl == 0) asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t_tile0));
        if (p.dbg & 16384) hs = 0u;                                  // (experiment: loop skeleton only)
        const uint32_t wc = __ballot_sync(0xFFFFFFFFu, hs != 0u);
        if (wc != 0u) {                                              // warp-uniform: a candidate in this warp
            // every load of the iteration is issued before the first one is waited for: the candidate's entry, its
            // own row, the tile's counter word and those of the earlier tiles of the group -- ONE round trip
            const int tt = tile - p.bd.tile_begin[sb];
            const int tg = tt & (kGrp - 1);
            uint4 ev = make_uint4(0u, 0u, 0u, 0u);
            Entry *e = p.tab + (hs - 1u);
            if (hs != 0u) ev = ld_entry(e);
            const unsigned long long w8 = __ldcg(p.l0 + tile);
            const unsigned long long g8 = wlane < tg ? __ldcg(p.l0 + tile - tg + wlane) : 0ull;
            float x[C];
            if (hs != 0u && !(p.dbg & 8192)) load_row<C, PERTURB>(p, idx, x);
            // candidates look at their entry again: still theirs -> first point of the cell, and the entry is zeroed
            // for the next call right away (a displaced candidate that reads the zero is not first either way)
            const bool f = hs != 0u && (ev.x | ev.y) != 0u && ev.x == (uint32_t)idx;
            if (f && !(p.dbg & 4096)) { if (p.dbg & 256) __stcg(reinterpret_cast<uint4 *>(e), make_uint4(0u, 0u, 0u, 0u)); else st_zero16_keep(e); }
            const uint32_t wf = __ballot_sync(0xFFFFFFFFu, f);
            if (wf != 0u) {                                          // warp-uniform: a first point in this warp
                // voxel index of the warp's first first point inside the sweep = first points of: earlier groups
                // (shared memory), earlier tiles of this group (one 8-byte load per lane), earlier warps of this
                // tile (its word)
                int pre_w;
                {
                    int s = tile_first_points(g8);
#pragma unroll
                    for (int d = 16; d > 0; d >>= 1) s += __shfl_xor_sync(0xFFFFFFFFu, s, d);
                    pre_w = s_E[p.bd.g1_begin[sb] + tt / kGrp] - s_base[sb] + s +
                            tile_first_points(w8 & ((1ull << (8 * (wid & 7))) - 1ull));
                }
                const int pre = pre_w + __popc(wf & ((1u << wlane) - 1u));
                const bool kept = f && pre < V && !(p.dbg & 2048);   // cells first seen after V voxels are refused
                const int voff = s_voff[sb];
                const uint32_t wk = __ballot_sync(0xFFFFFFFFu, kept);
                // the cell's later points: min(cnt, P-1) arrivals (slot0 in the entry, the rest in its row) and, for
                // a cell with more than P points, the levels of its overflow list
                const uint32_t D = (uint32_t)P - 1u;
                const uint32_t m = f ? (ev.z < D ? ev.z : D) : 0u;
                uint32_t *row = p.lists + (size_t)(hs - 1u) * LS;    // (only dereferenced by first points)
                uint32_t *L = nullptr;
                if (f && P > 1 && ev.z >= D) {
                    const uint32_t lp = __ldcg(row + D - 1u);
                    if (lp != 0u) { L = p.ovf + (size_t)(lp - 1u) * LS; __stcg(row + D - 1u, 0u); }
                }
                if (wk != 0u) {
                    // the warp's kept voxels are the rows [voff + pre_w, + popc(wk)): zero them with coalesced stores,
                    // then (after __syncwarp, which orders the stores of the participating lanes) every first point
                    // fills its row
                    warp_zero_fill(p.voxels + (size_t)(voff + pre_w) * PC, __popc(wk) * PC, wlane);
                    __syncwarp();
                }
                if (kept) {
                    const long long v = voff + pre;
                    float *vrow = p.voxels + v * PC;
                    float sum[C];
                    if constexpr (C == 4) {
                        reinterpret_cast<float4 *>(vrow)[0] = make_float4(x[0], x[1], x[2], x[3]);
                    } else {
#pragma unroll
                        for (int c = 0; c < C; ++c) vrow[c] = x[c];
                    }
#pragma unroll
                    for (int c = 0; c < C; ++c) sum[c] = x[c];       // 0 + x == x: the oracle's sum starts at 0.0f
                    const int cx = axis_cell(x[0], p.gd, 0), cy = axis_cell(x[1], p.gd, 1), cz = axis_cell(x[2], p.gd, 2);
                    if (p.p2v) p.p2v[idx] = (int)v;
                    if (p.p2s) p.p2s[idx] = 0;
                    int n = 1;
                    if (m > 0u && L == nullptr && D <= (uint32_t)kFastD) {
                        // the usual cell (at most P points): its arrivals -- slot0 from the entry, the others with
                        // independent loads from its row -- are sorted by point order in shared memory, then gathered
                        // with the next point's row in flight while the current one is stored
                        uint32_t *sc = s_sc + tid * p.DS;
                        sc[0] = ev.w;
                        for (uint32_t j = 0; j + 1u < m; ++j) sc[j + 1u] = __ldcg(row + j);
                        for (uint32_t a = 1; a < m; ++a) {           // insertion sort, largest W (earliest point) first
                            const uint32_t t = sc[a];
                            uint32_t b = a;
                            while (b > 0u && sc[b - 1u] < t) { sc[b] = sc[b - 1u]; --b; }
                            sc[b] = t;
                        }
                        float xn[C];
                        load_row<C, PERTURB>(p, (int)(kWBase - sc[0]), xn);
                        for (uint32_t q = 1; q <= m; ++q) {
                            const int i2 = (int)(kWBase - sc[q - 1u]);
#pragma unroll
                            for (int c = 0; c < C; ++c) x[c] = xn[c];
                            if (q < m) load_row<C, PERTURB>(p, (int)(kWBase - sc[q]), xn);
                            if constexpr (C == 4) {
                                reinterpret_cast<float4 *>(vrow)[q] = make_float4(x[0], x[1], x[2], x[3]);
                            } else {
#pragma unroll
                                for (int c = 0; c < C; ++c) vrow[q * C + c] = x[c];
                            }
#pragma unroll
                            for (int c = 0; c < C; ++c) sum[c] = __fadd_rn(sum[c], x[c]);
                            if (p.p2v) p.p2v[i2] = (int)v;
                            if (p.p2s) p.p2s[i2] = (int)q;
                        }
                        n = 1 + (int)m;
                    } else if (m > 0u) {
                        // crowded cell (overflow list) or a very large P: selection straight from global memory --
                        // slot q takes the largest W below slot q-1's among arrivals and overflow levels
                        uint32_t prev = 0xFFFFFFFFu;
                        for (int q = 1; q < P; ++q) {
                            uint32_t best = 0u;
                            if (ev.w < prev) best = ev.w;
                            for (uint32_t j = 0; j + 1u < m; ++j) {
                                const uint32_t t = __ldcg(row + j);
                                if (t < prev && t > best) best = t;
                            }
                            if (L) {
                                for (uint32_t j = 0; j < D; ++j) {   // (sorted: the first one below prev is the largest)
                                    const uint32_t t = __ldcg(L + j);
                                    if (t < prev) { if (t > best) best = t; break; }
                                }
                            }
                            if (best == 0u) break;
                            prev = best;
                            const int i2 = (int)(kWBase - best);
                            load_row<C, PERTURB>(p, i2, x);
                            if constexpr (C == 4) {
                                reinterpret_cast<float4 *>(vrow)[q] = make_float4(x[0], x[1], x[2], x[3]);
                            } else {
#pragma unroll
                                for (int c = 0; c < C; ++c) vrow[q * C + c] = x[c];
                            }
#pragma unroll
                            for (int c = 0; c < C; ++c) sum[c] = __fadd_rn(sum[c], x[c]);
                            if (p.p2v) p.p2v[i2] = (int)v;
                            if (p.p2s) p.p2s[i2] = q;
                            n = q + 1;
                        }
                    }
                    p.num_points[v] = n;
                    reinterpret_cast<int4 *>(p.coors)[v] = make_int4(sb, cz, cy, cx);
                    if (p.feats) {
                        const float fn = (float)n;
#pragma unroll
                        for (int c = 0; c < C; ++c)
                            if (c < F) p.feats[v * F + c] = __fdiv_rn(sum[c], fn);
                    }
                }
                // kept or refused, the cell's first point leaves the row and the overflow list zero for the next call
                if (f) {
                    for (uint32_t j = 0; j + 1u < m; ++j) st_zero4_keep(row + j);
                    if (L) {
                        for (uint32_t j = 0; j < D; ++j) {
                            if (__ldcg(L + j) == 0u) break;
                            __stcg(L + j, 0u);
                        }
                    }
                }
            }
        }
        if (p.stamps && (p.dbg & 64) && tl == 0 && tile < 5000) {    // per-tile record of phase B (warp 0 of the team)
            unsigned long long t1;
            asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t1));
            unsigned long long *rec = p.stamps + 16 * 296 + 3 * (size_t)(5000 + tile);
            rec[0] = t_tile0; rec[1] = t1; rec[2] = ((unsigned long long)blockIdx.x << 32) | (unsigned)__popc(wc);
        }
        hs = hs_nxt;
    }
    if (p.stamps && (p.dbg & 65536) && wlane == 0 && blockIdx.x < 64) {   // (experiment) when every warp leaves the emit loop
        unsigned long long t1;
        asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t1));
        p.stamps[16 * 296 + 3 * 10000 + blockIdx.x * 32 + wid] = t1;
    }
    if (p.stamps) __syncthreads();
    stamp(p, 5);
    // The last CTA to get here zeroes the counters and signs the workspace as clean.  (No fence: what the others
    // READ from the counters has arrived -- they used it -- before their thread 0 counts them in, and everything
    // written here only has to be visible to the next launch.)
    __syncthreads();
    if (tid == 0) s_flag = (atomicAdd(&p.hdr->done, 1u) == (unsigned)(nct - 1)) ? 1 : 0;
    __syncthreads();
    if (s_flag) {
        for (int j = tid; j < p.ntiles; j += kPT) p.l0[j] = 0ull;
        for (int j = tid; j < p.ng1; j += kPT) p.l1[(size_t)j * kL1Stride] = 0;
        if (tid == 0) {
            unsigned long long chi = p.front_end;
            if (was_clean) chi = __ldcg(&p.hdr->clean_hi);
            p.hdr->clean_hi = chi < p.back_off ? chi : p.back_off;
            p.hdr->ws_bytes = p.ws_bytes;
            p.hdr->done = 0u;
            p.hdr->next_a = 0u;
            p.hdr->next_b = 0u;
            p.hdr->magic = kWsMagic;
        }
    }
    stamp(p, 6);
}

}  // namespace rpc
