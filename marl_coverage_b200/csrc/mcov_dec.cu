// mcov_dec.cu — DecGridRL reset/step/observe as ONE fused sm_100a kernel (auto-reset and episode
// statistics included: a step is a single launch).
//
// Reference semantics restated here (file:line in ExistentialRobotics/MARL-Coverage):
//   DecGridRL.step            Environments/dec_grid_rl.py:91-169
//   updateRobotPos/isOccupied Environments/dec_grid_rl.py:171-204,297-310
//   observe                   Environments/dec_grid_rl.py:206-258
//   SquareSensor              Environments/Sensors/squaresensor.py:15-37
//   LidarSensor               Environments/Sensors/lidar.py:16-65
//   get_egocentric_observations / arraySubset   dec_grid_rl.py:312-372,393-421
//   updateCommmunicationGraph dec_grid_rl.py:374-391
//   reset                     dec_grid_rl.py:449-531
//   done / percent_covered    dec_grid_rl.py:533-552
//
// Mapping: a GROUP of G lanes (8, 16 or 32 >= numrobot) owns one env per loop iteration, so a warp
// steps 32/G envs at once (persistent group-stride loop, 256 threads per CTA, 4 CTAs per SM -> up
// to 128 envs in flight per SM, no CTA-wide barrier on the path).  Lanes < N of a group hold the
// robots' poses for the sequential collision-checked moves (shuffle + ballot); afterwards lanes own
// (robot, window-row) pairs.  All per-robot maps are bit planes in HBM ("strip-major", see
// include/mcov.h); a window row is a <=32-bit field cut from <=2 words with a funnel shift.  The
// words under the observation window are loaded ONCE and serve both the memory update
// (free_i/obst_i |= sensor hits) and the observation rows.  Newly covered cells are counted with
// __popc + warp reductions, no global atomics: every bitmap word has exactly one writer (owner
// election when footprints overlap).  Observation rows are ORed into a per-env bit stream in
// shared memory (bit k = float k of the env's obs block); the float32 block is expanded from it
// (4 bits -> one float4 through a 16-entry shared-memory table) and stored with coalesced 128-bit
// streaming stores straight into the torch buffer.  The next env's header, poses
// and actions are prefetched into registers while the current env is processed.
#include <stdlib.h>

#include "mcov_internal.h"
#include "mcov_dec_shared.cuh"

namespace mcov {

namespace {

constexpr int MAX_WARPS = 8;
#ifndef MCOV_DEC_MINB
#define MCOV_DEC_MINB 4                      // resident 256-thread CTAs per SM the register budget is set for
#endif

// ---- lane groups: G consecutive lanes (8, 16 or 32) own one env; a warp steps 32/G envs at once.
// Every warp-level primitive below names only the group's lanes, so groups of one warp may diverge
// (e.g. one of them re-seeding a finished env) without deadlock.
template <int G>
struct Grp {
    int lane, lg, gbase;
    unsigned mask;
    __device__ __forceinline__ explicit Grp(int lane_) : lane(lane_), lg(lane_ & (G - 1)), gbase(lane_ & ~(G - 1)) {
        mask = G == 32 ? FULL : (((1u << (G & 31)) - 1u) << gbase);
    }
    __device__ __forceinline__ int shfl(int v, int src) const { return __shfl_sync(mask, v, src, G); }
    __device__ __forceinline__ uint32_t ballot(bool pred) const {
        return (__ballot_sync(mask, pred) >> gbase) & (G == 32 ? FULL : ((1u << (G & 31)) - 1u));
    }
    __device__ __forceinline__ bool any(bool pred) const { return __any_sync(mask, pred) != 0; }
    __device__ __forceinline__ int sum(int v) const { return (int)__reduce_add_sync(mask, v); }
    __device__ __forceinline__ void sync() const { __syncwarp(mask); }
};

// Raw (untouched) register copies of the next env's inputs: nothing is computed on them until the
// env is processed, so the loads stay in flight behind the current env's work.
struct Prefetch {
    uint32_t pos;        // lg < N: packed pose
    uint32_t act8;       // lg < N: per-robot move (U8 format)
    long long act64;     // base-4 joint action (I64 format)
    int map_idx;
    uint32_t shown;
};

__device__ __forceinline__ Prefetch load_prefetch(const DecParams& p, int e, int lg) {
    Prefetch f;
    const unsigned char* hrec = p.hdr + (size_t)e * p.hdr_stride;
    const EnvHead* head = reinterpret_cast<const EnvHead*>(hrec);
    f.map_idx = head->map_idx;
    f.shown = head->flags;
    f.pos = 0;
    f.act8 = 0;
    f.act64 = 0;
    if (lg < p.N) f.pos = reinterpret_cast<const uint32_t*>(hrec + p.off_pos)[lg];
    if (p.mode == MODE_STEP) {
        if (p.act_fmt == MCOV_ACT_I64_BASE4) f.act64 = reinterpret_cast<const long long*>(p.actions)[e];
        else if (lg < p.N) f.act8 = reinterpret_cast<const uint8_t*>(p.actions)[(size_t)e * p.N + lg];
    }
    return f;
}

// per-group shared memory
struct GroupSmem {
    double* stats;       // [MCOV_STATS_LEN] running sums of this group (lg 0 only)
    uint32_t* bits;      // [bits_words] observation bit stream: bit k = float k of the env's obs block
    uint32_t* posF;      // [N] poses after the moves
    uint32_t* nearM;     // [N] robots whose sensor rows may share a bitmap word with robot i
    int* covS;           // [N] per-robot newly known-free cells of this step
    uint32_t* winF;      // lidar only: [N*S] sensor-window free hits, window-relative bits
    uint32_t* winO;      //             [N*S] obstacle hits
    uint32_t* wallW;     //             [N*S] truth wall window
    const uint32_t* mapW;  // Spec::SMT: the shared-memory copy of the truth-map table (wall planes, walkable planes)
    const uint32_t* mapP;
};

__device__ __forceinline__ GroupSmem carve(unsigned char* base, const DecParams& p) {
    GroupSmem s;
    s.stats = reinterpret_cast<double*>(base);
    s.bits = reinterpret_cast<uint32_t*>(base + MCOV_STATS_LEN * 8);
    s.posF = s.bits + (p.st_slot_words ? 2 * p.st_slot_words : p.bits_words);   // Spec::ST: two bit-stream slots
    s.nearM = s.posF + p.N;
    s.covS = reinterpret_cast<int*>(s.nearM + p.N);
    s.winF = reinterpret_cast<uint32_t*>(s.covS + p.N);
    s.winO = s.winF + p.N * (p.S + 1);                 // S window rows + the dummy row of the beam table walk
    s.wallW = s.winO + p.N * (p.S + 1);
    return s;
}

#ifndef MCOV_ST_UNROLL
#define MCOV_ST_UNROLL 4                     // unroll of the observation store loop
#endif
#define MCOV_PRAGMA_(x) _Pragma(#x)
#define MCOV_UNROLL(n) MCOV_PRAGMA_(unroll n)
#ifndef MCOV_LUT8
#define MCOV_LUT8 0                          // observation store: byte -> 2 x float4 table (8 kB of shared memory per CTA): 1 = two 128-bit stores, 2 = one 256-bit store
#endif
#ifndef MCOV_PF_TRIM
#define MCOV_PF_TRIM 0                       // L1 prefetch before the band-row loop: 0 first+last row of every run, 1 one per 1-bit run, 2 own planes only
#endif
#ifndef MCOV_PF_EARLY
#define MCOV_PF_EARLY 0                      // 1: prefetch the window runs for the pre-move band (grown by one cell) before the moves
#endif
#ifndef MCOV_NO_PF
#define MCOV_NO_PF 0                         // 1: drop the L1 prefetch of the window runs before the band-row loop
#endif
#ifndef MCOV_LIDAR_RUNTAB
#define MCOV_LIDAR_RUNTAB 0                  // 1: beam table walk over host-built same-row runs — parity-green, A/B on C3: 0.129 vs 0.104 ms
                                             // cell by cell (a warp's beams have 1..10 runs: the longest decides, and a run costs more than a cell)
#endif
#ifndef MCOV_LIDAR_PERM
#define MCOV_LIDAR_PERM 0                    // beam table walk: spread the beams a warp walks at once over the circle (A/B on C3: 0.1019 vs 0.1022 ms, noise; off)
#endif
#ifndef MCOV_LIDAR_RUNS
#define MCOV_LIDAR_RUNS 0                    // beam table walk: one shared-memory atomic per same-row run of cells (A/B on C3: 0.1108 vs 0.1022 ms, slower; off)
#endif
#ifndef MCOV_PF_NEXT
#define MCOV_PF_NEXT 0                       // measured on B200 (C5): 0.213 -> 0.233 ms with it on — the step is DRAM-throughput bound, extra line fetches cost more than the latency they hide
#endif

// One pass over env e in `mode`.  Returns true when a finished env must be re-seeded in place by a
// second (MODE_RESET) pass of the same lane group (fused auto-reset).
// Compile-time shape of an env (0 = read the value from DecParams at run time).  The named
// BASELINE configs get their own instantiation so that loop bounds, divisions and masks fold.
template <int G_, bool LIDAR_, int CN_, int CR_, int CE_, bool ST_ = false, bool SMT_ = false>
struct Spec {
    static constexpr int G = G_, CN = CN_, CR = CR_, CE = CE_;
    static constexpr bool LIDAR = LIDAR_;
    static constexpr bool FIXED = CN_ > 0;             // then C == 3, N, R, ego are constants
    static constexpr bool ST = ST_;                    // observation blocks leave through the CTA's store warp (TMA bulk stores)
    // truth-map table resident in shared memory: one 1024-thread CTA per SM stages the whole read-only table (wall +
    // walkable planes of every map) once per launch with TMA bulk copies; the sensor reads it with shared-memory loads
    static constexpr bool SMT = SMT_;
    static constexpr int THREADS = SMT_ ? 1024 : 256, MINB = SMT_ ? 1 : MCOV_DEC_MINB;
};

// a word of the truth-map table: read-only global memory, or the shared-memory copy (Spec::SMT)
template <bool SMT>
__device__ __forceinline__ uint32_t ld_map(const uint32_t* ptr) { return SMT ? *ptr : __ldg(ptr); }

// ---- store-warp hand-over (Spec::ST) -------------------------------------------------------------
// The lane groups of the pass warps do not write their observation block themselves: a group fills one of
// its two bit-stream slots and signals the slot's `full` mbarrier; the CTA's last warp (the store warp)
// expands a full slot into a float32 staging buffer in shared memory, frees the slot (`empty` mbarrier) and
// hands the whole block to the TMA engine (cp.async.bulk shared -> global, one 5.8 kB burst per env at C2).
// Pass warps therefore never wait for store issue, and DRAM sees whole-block bursts from 4 store warps per
// SM instead of 128 lane groups each trickling 16-byte stores (tools/ubench/store_patterns.cu: 4.65 -> 6.0 TB/s).
struct StoreLink {
    uint32_t pub, con;        // shared-memory addresses of this group's counters: blocks published / consumed
    int it;                   // blocks published so far by this group (slot = it & 1)
};
__device__ __forceinline__ uint32_t ld_acquire_smem(uint32_t addr) {
    uint32_t v;
    asm volatile("ld.acquire.cta.shared.u32 %0, [%1];" : "=r"(v) : "r"(addr) : "memory");
    return v;
}
__device__ __forceinline__ void st_release_smem(uint32_t addr, uint32_t v) {
    asm volatile("st.release.cta.shared.u32 [%0], %1;" :: "r"(addr), "r"(v) : "memory");
}
// slot it & 1 was last used by block it - 2: wait until the store warp has consumed it
__device__ __forceinline__ void wait_slot_free(const StoreLink& link) {
    while ((int)ld_acquire_smem(link.con) < link.it - 1) __nanosleep(32);
}

// While this env's observation block streams out, start the DRAM reads the group's NEXT env will need:
// the lines under every robot's band in its own {free, obst} planes, the union plane and the two truth
// planes.  Poses are the pre-move ones (a move shifts a window by one cell, i.e. almost always inside the
// same 128-byte lines), so this is a hint, never a dependency.
template <class SP>
__device__ __forceinline__ void prefetch_next_env(const DecParams& p, const Prefetch& nx, const int en,
                                                  const Grp<SP::G>& g) {
    constexpr int G = SP::G;
    const int N = SP::FIXED ? SP::CN : p.N, R = SP::FIXED ? SP::CR : p.R, ego = SP::FIXED ? SP::CE : p.ego;
    const int rad = ego > R ? ego : R, RW = 2 * rad + 1;
    const int W = p.g.W, Wp = p.g.Wp, strips = p.g.strips, PW = p.g.PW;
    const uint32_t eN = (uint32_t)en * (uint32_t)N, vis_off = (uint32_t)en * (uint32_t)PW;
    const uint32_t map_off = (uint32_t)nx.map_idx * (uint32_t)PW;
    for (int base = 0; base < 2 * N; base += G) {
        const int idx = base + g.lg, i = idx >> 1, k = idx & 1;
        const uint32_t pi = (uint32_t)g.shfl((int)nx.pos, i & (G - 1));
        if (idx >= 2 * N) continue;
        const int xi = unpack_x(pi), yi = unpack_y(pi);
        const int s = ((yi - rad) >> 5) + k;
        if (s < 0 || s >= strips || (k == 1 && ((yi - rad) & 31) + RW < 32)) continue;
        const int xa = max(xi - rad - 1, 0), xb = min(xi + rad + 1, W - 1);
        const uint32_t fw = (eN + i) * PW + s * Wp;
        prefetch_l2(p.fo + fw + xa);
        prefetch_l2(p.fo + fw + xb);
        if (s >= ((yi - R - 1) >> 5) && s <= ((yi + R + 1) >> 5)) {
            const int xc = max(xi - R - 1, 0), xd = min(xi + R + 1, W - 1);
            prefetch_l2(p.visited + vis_off + s * Wp + xc);
            prefetch_l2(p.visited + vis_off + s * Wp + xd);
            prefetch_l2(p.wall + map_off + s * Wp + xc);
            prefetch_l2(p.wall + map_off + s * Wp + xd);
            if (!SP::LIDAR) {
                prefetch_l2(p.pos + map_off + s * Wp + xc);
                prefetch_l2(p.pos + map_off + s * Wp + xd);
            }
        }
    }
}


template <class SP>
__device__ __forceinline__ int env_pass(const DecParams& p, const GroupSmem& sm, const float4* lut, const float4* lut8,
                                         const int e, const int mode, const Prefetch& cur,
                                         const Grp<SP::G>& g, const Prefetch& nxt, const int e_next,
                                         const StoreLink& link) {
    constexpr int G = SP::G;
    constexpr bool LIDAR = SP::LIDAR;
    const int N = SP::FIXED ? SP::CN : p.N, R = SP::FIXED ? SP::CR : p.R, ego = SP::FIXED ? SP::CE : p.ego;
    const int C = SP::FIXED ? 3 : p.C;
    // the compile-time shapes are only dispatched with these options off (launch_dec), so they fold away
    const bool single_tool = !SP::FIXED && p.single_tool;
    const int comm_radius = SP::FIXED ? 0 : p.comm_radius;
    const int phases = SP::FIXED ? (PH_MOVE | PH_SENSE) : p.phases;          // split only for dist_reward (C == 4)
    const int S = 2 * R + 1, Sp = S + 1, ow = 2 * ego + 1, rad = ego > R ? ego : R, RW = 2 * rad + 1;
    const int W = p.g.W, L = p.g.L, Wp = p.g.Wp, strips = p.g.strips, PW = p.g.PW;
    const int lg = g.lg;
    auto div_n = [&](int k) { return SP::FIXED ? k / (SP::FIXED ? SP::CN : 1) : fast_div(k, p.n_magic); };
    auto div_s = [&](int k) { return SP::FIXED ? k / (2 * SP::CR + 1) : fast_div(k, p.s_magic); };
    auto div_rw = [&](int k) {
        return SP::FIXED ? k / (2 * (SP::CE > SP::CR ? SP::CE : SP::CR) + 1) : fast_div(k, p.rw_magic);
    };

    unsigned char* hrec = p.hdr + (size_t)e * p.hdr_stride;
    EnvHead* head = reinterpret_cast<EnvHead*>(hrec);
    uint32_t* hpos = reinterpret_cast<uint32_t*>(hrec + p.off_pos);
    int32_t* hcov = reinterpret_cast<int32_t*>(hrec + p.off_cov);
    // 32-bit word offsets from the arena bases (launch_dec checks E*N*PW < 2^32): one IMAD.WIDE per
    // access instead of carrying / rebuilding 64-bit per-env pointers
    const uint32_t eN = (uint32_t)e * (uint32_t)N;
    const uint32_t vis_off = (uint32_t)e * (uint32_t)PW;
    uint2* const fo_g = p.fo;
    uint32_t* const vis_g = p.visited;

    g.sync();                                    // previous pass has finished reading sm.bits
    const bool fresh = (mode == MODE_RESET);
    int map_idx;
    uint32_t shown, myv;
    if (fresh) {
        int mi = 0;
        if (lg == 0) draw_reset(p, e, head->episode, sm.posF, &mi);
        map_idx = g.shfl(mi, 0);
        if (!p.skip_wipe) {
            // wipe this env's planes (dec_grid_rl.py:505-514) with 128-bit stores
            uint4* z = reinterpret_cast<uint4*>(fo_g + (size_t)eN * PW);
            const int n16 = N * (PW / 2);
            for (int k = lg; k < n16; k += G) z[k] = make_uint4(0, 0, 0, 0);
            uint4* zv = reinterpret_cast<uint4*>(vis_g + vis_off);
            for (int k = lg; k < PW / 4; k += G) zv[k] = make_uint4(0, 0, 0, 0);
        }
        g.sync();
        myv = lg < N ? sm.posF[lg] : 0u;
        shown = 0;
    } else {
        map_idx = cur.map_idx;
        shown = cur.shown;
        myv = cur.pos;
    }
    const uint32_t map_off = (uint32_t)map_idx * (uint32_t)PW;
    const uint32_t* const wall_g = SP::SMT ? sm.mapW : p.wall;
    const uint32_t* const tpos_g = SP::SMT ? sm.mapP : p.pos;

    // early loads for the bookkeeping at the end of the pass (kept in flight until then)
    int cs0 = 0, fc = 1;
    double th = 0.0, epr0 = 0.0;
    if (mode == MODE_STEP && lg == 0) {
        cs0 = head->currstep;
        th = head->done_thresh;
        epr0 = head->ep_return;
        fc = __ldg(p.free_counts + map_idx);
    }
    int cov_old = 0;
    if (mode == MODE_STEP && lg < N) cov_old = hcov[lg];

#if MCOV_PF_EARLY
    if (!fresh && mode == MODE_STEP) {
        // The band can only move by one cell: pull the lines under the PRE-move band, grown by one row either side and
        // one cell up and down, towards the SM before the wall bits are fetched and the moves resolved — the loads of
        // the band-row loop are then a move's worth of instructions behind their prefetch instead of a few dozen.
        for (int base = 0; base < 2 * N; base += G) {             // every lane of the group takes part in the shuffle
            const int idx = base + lg, i = idx >> 1, k = idx & 1;
            const uint32_t pi = (uint32_t)g.shfl((int)myv, i & (G - 1));
            if (idx >= 2 * N) continue;
            const int xi = unpack_x(pi), yi = unpack_y(pi);
            const int yb = yi - rad - 1;
            const int s = (yb >> 5) + k;
            if (s < 0 || s >= strips || (k == 1 && (yb & 31) + RW + 2 <= 32)) continue;
            const int xa = max(xi - rad - 1, 0), xb = min(xi + rad + 1, W - 1);
            const uint32_t fw = (eN + i) * PW + s * Wp;
            prefetch_l1(fo_g + fw + xa);
            prefetch_l1(fo_g + fw + xb);
            if (s >= ((yi - R - 1) >> 5) && s <= ((yi + R + 1) >> 5)) {
                const int xc = max(xi - R - 1, 0), xd = min(xi + R + 1, W - 1);
                prefetch_l1(vis_g + vis_off + s * Wp + xc);
                prefetch_l1(vis_g + vis_off + s * Wp + xd);
                prefetch_l1(wall_g + map_off + s * Wp + xc);
                prefetch_l1(wall_g + map_off + s * Wp + xd);
                if (!LIDAR) {
                    prefetch_l1(tpos_g + map_off + s * Wp + xc);
                    prefetch_l1(tpos_g + map_off + s * Wp + xd);
                }
            }
        }
    }
#endif
    int px = unpack_x(myv), py = unpack_y(myv);
    int ncoll = 0;
    bool sentinel = false;
    if (mode == MODE_STEP) {
        // ---- action decode (dec_grid_rl.py:110-117) ------------------------------------------
        int u;
        if (p.act_fmt == MCOV_ACT_I64_BASE4) {
            sentinel = (cur.act64 == -1);
            u = (int)((cur.act64 >> (2 * lg)) & 3);
        } else {
            u = (int)(cur.act8 & 3u);
        }
        if (!sentinel && (phases & PH_MOVE)) {
            // ---- sequential, collision-checked moves (dec_grid_rl.py:128-145,171-204) --------
            // Dec table: 0:x+1  1:y+1  2:x-1  3:y-1
            const int tx = px + (u == 0) - (u == 2);
            const int ty = py + (u == 1) - (u == 3);
            int blocked0 = 1;
            if (lg < N && tx >= 0 && tx < W && ty >= 0 && ty < L) blocked0 = (ld_map<SP::SMT>(wall_g + map_off + (ty >> 5) * Wp + tx) >> (ty & 31)) & 1u;
            const uint32_t tv = pack_xy(tx, ty);
            uint32_t pv = myv;
            for (int k = 0; k < N; ++k) {
                const uint32_t tvk = (uint32_t)g.shfl((int)tv, k);
                const int blk = g.shfl(blocked0, k);
                const uint32_t b = g.ballot((lg < N) && pv == tvk);
                if (!blk && b == 0) {
                    if (lg == k) pv = tvk;
                    shown |= 1u << k;
                } else {
                    ++ncoll;
                }
            }
            px = unpack_x(pv);
            py = unpack_y(pv);
        } else if (!sentinel) {
            // moves were resolved by an earlier PH_MOVE launch (dist_reward pipeline)
            const uint2 ms = p.move_scratch[e];
            ncoll = (int)ms.x;
            shown = ms.y;
        }
    }
    // publish poses for cross-robot lookups and persist them; clear the per-pass scratch
    const uint32_t newv = pack_xy(px, py);
    if (lg < N) {
        sm.posF[lg] = newv;
        sm.covS[lg] = 0;
        sm.nearM[lg] = 0;
        if (mode != MODE_OBSERVE && (fresh || newv != myv)) hpos[lg] = newv;
    }
    if (SP::ST) wait_slot_free(link);          // the store warp has finished reading this slot
    for (int k = lg; k < p.bits_words; k += G) sm.bits[k] = 0;

    // ---- communication graph (dec_grid_rl.py:374-391): Chebyshev ball, diagonal included ------
    if (mode != MODE_OBSERVE && (comm_radius > 0 || fresh)) {
        uint32_t myadj = 1u << lg;
        if (comm_radius > 0) {
            for (int i = 0; i < N; ++i) {
                const int xi = g.shfl(px, i), yi = g.shfl(py, i);
                const uint32_t b = g.ballot((lg < N) && max(abs(px - xi), abs(py - yi)) <= comm_radius);
                if (lg == i) myadj = b;
            }
        }
        if (lg < N) reinterpret_cast<uint32_t*>(hrec + p.off_adj)[lg] = myadj;
    }
    if (mode == MODE_STEP && !(phases & PH_SENSE)) {
        if (lg == 0) p.move_scratch[e] = make_uint2((uint32_t)ncoll, shown);
        return PASS_DONE;
    }
    g.sync();

    const bool do_sense = fresh || (mode == MODE_STEP && !sentinel);
    if (!fresh && !MCOV_NO_PF && !(MCOV_PF_EARLY && mode == MODE_STEP)) {
        // Pull the lines under every robot's window towards the SM now: band rows are contiguous in
        // x within a strip ((2*rad+1) * 8 B <= 2 lines), so touching the first and last row of each
        // run covers it.  One lane per (robot, strip); the slot loop below then hits L1/L2 instead
        // of paying one DRAM round trip per iteration.
        for (int idx = lg; idx < 2 * N; idx += G) {
            const int i = idx >> 1, k = idx & 1;
            const uint32_t pi = sm.posF[i];
            const int xi = unpack_x(pi), yi = unpack_y(pi);
            const int s = ((yi - rad) >> 5) + k;
            if (s < 0 || s >= strips || (k == 1 && ((yi - rad) & 31) + RW <= 32)) continue;
            const int xa = max(xi - rad, 0), xb = min(xi + rad, W - 1);
            const uint32_t fw = (eN + i) * PW + s * Wp;
            prefetch_l1(fo_g + fw + xa);
            prefetch_l1(fo_g + fw + xb);
#if MCOV_PF_TRIM == 2
            // only the robot's own {free, obst} runs (the state that comes from DRAM); the union plane and the
            // truth planes are left to the band-row loop
#elif MCOV_PF_TRIM == 1
            // the sensor rows of a 1-bit plane are (2R+1) * 4 B: one prefetch per run, the rare run that crosses a
            // 128-byte line takes its miss in the band-row loop
            if (do_sense && s >= ((yi - R) >> 5) && s <= ((yi + R) >> 5)) {
                prefetch_l1(vis_g + vis_off + s * Wp + xi);
                prefetch_l1(wall_g + map_off + s * Wp + xi);
                if (!LIDAR) prefetch_l1(tpos_g + map_off + s * Wp + xi);
            }
#else
            if (do_sense && s >= ((yi - R) >> 5) && s <= ((yi + R) >> 5)) {
                const int xc = max(xi - R, 0), xd = min(xi + R, W - 1);
                prefetch_l1(vis_g + vis_off + s * Wp + xc);
                prefetch_l1(vis_g + vis_off + s * Wp + xd);
                if (SP::SMT) continue;                             // the truth table is in shared memory
                prefetch_l1(wall_g + map_off + s * Wp + xc);
                prefetch_l1(wall_g + map_off + s * Wp + xd);
                if (!LIDAR) {
                    prefetch_l1(tpos_g + map_off + s * Wp + xc);
                    prefetch_l1(tpos_g + map_off + s * Wp + xd);
                }
            }
#endif
        }
    }
    if (do_sense) {
        // which robots' sensor rows can share a bitmap word with robot i's?  (one lane per pair)
        for (int idx = lg; idx < N * N; idx += G) {
            const int i = div_n(idx), j = idx - i * N;
            const uint32_t pi = sm.posF[i], pj = sm.posF[j];
            const int yi = unpack_y(pi), yj = unpack_y(pj);
            const bool nr = (i != j) && abs(unpack_x(pi) - unpack_x(pj)) <= 2 * R &&
                            max((yi - R) >> 5, (yj - R) >> 5) <= min((yi + R) >> 5, (yj + R) >> 5);
            if (nr) atomicOr(&sm.nearM[i], 1u << j);
        }
        if (LIDAR) {
            // lidar.py:16-65 — stage the truth wall window, then march beams in float64
            for (int idx = lg; idx < N * S; idx += G) {
                const int i = div_s(idx), r = idx - i * S;
                const uint32_t pi = sm.posF[i];
                const int x = unpack_x(pi) - R + r, y0 = unpack_y(pi) - R;
                const uint32_t wrow = (x >= 0 && x < W) ? read_window(wall_g + map_off, Wp, strips, x, y0, S) : 0xffffffffu;
                sm.wallW[i * Sp + r] = wrow;
                // table walk: every beam starts on the robot's own cell, lidar.py:30 — set it here once instead of
                // by n_lasers colliding shared-memory atomics.  (A robot never stands on an obstacle inside the env;
                // a sensor probe can put one there: then every beam ends at once on that cell, lidar.py:58-63.)
                const bool own_wall = (wrow >> R) & 1u;
                sm.winF[i * Sp + r] = (p.lidar_cells && r == R && !own_wall) ? (1u << R) : 0u;
                sm.winO[i * Sp + r] = (p.lidar_cells && r == R && own_wall) ? (1u << R) : 0u;
            }
            if (lg < N) sm.wallW[lg * Sp + S] = 0xffffffffu;        // dummy row of the beam table walk: all wall
            g.sync();
            if (p.lidar_cells) {
                // integer beam table (proved start-independent on the host, mcov_set_lidar_beams):
                // walk the <= 16 window cells of the beam until the first wall
                const int nl = p.n_lasers;
                for (int idx = lg; idx < N * nl; idx += G) {
                    // neighbouring lanes take the same beam of different robots: beams of one robot share
                    // cells near the origin and would collide on the same shared-memory words
                    const int l0 = div_n(idx), i = idx - l0 * N;
#if MCOV_LIDAR_PERM
                    // ... and the 32/N beams a warp walks at once are spread over the circle (l -> l * m mod n_lasers,
                    // m ~ n_lasers / 4 coprime to it): adjacent beams run through the same cells for their first steps
                    const int lm = l0 * (int)p.l_mult;
                    const int l = lm - nl * fast_div(lm, p.l_magic);
#else
                    const int l = l0;
#endif
                    const uint32_t* ww = sm.wallW + i * Sp;
                    uint32_t* wf = sm.winF + i * Sp;
                    uint32_t* wo = sm.winO + i * Sp;
                    if ((ww[R] >> R) & 1u) continue;               // own cell is an obstacle: the beam ended there
#if MCOV_LIDAR_RUNTAB
                    walk_beam_runs(p.lidar_runs, l, ww, wf, wo);
#else
                    walk_beam_cells(p.lidar_cells, l, ww, wf, wo);
#endif
                }
            } else {
            const double rng = (double)R;
            for (int idx = lg; idx < N * p.n_lasers; idx += G) {
                const int i = idx / p.n_lasers, l = idx - i * p.n_lasers;
                const uint32_t pi = sm.posF[i];
                const int xi = unpack_x(pi), yi = unpack_y(pi);
                const double xinc = __ldg(p.beams + 3 * l), yinc = __ldg(p.beams + 3 * l + 1),
                             dinc = __ldg(p.beams + 3 * l + 2);
                double cx = (double)xi, cy = (double)yi, dist = 0.0;
                int rx, ry;
                bool walk;
                while (true) {
                    const bool inside = cx >= 0.0 && cy >= 0.0 && cx < (double)W && cy < (double)L;
                    rx = __double2int_rz(cx) - (xi - R);
                    ry = __double2int_rz(cy) - (yi - R);
                    const bool inwin = rx >= 0 && rx < S && ry >= 0 && ry < S;
                    walk = inside && inwin && !((sm.wallW[i * Sp + rx] >> ry) & 1u);
                    if (!(walk && dist < rng)) break;
                    atomicOr(&sm.winF[i * Sp + rx], 1u << ry);
                    cx += xinc;
                    cy += yinc;
                    dist += dinc;
                }
                if (rx >= 0 && rx < S && ry >= 0 && ry < S) {
                    if (walk) atomicOr(&sm.winF[i * Sp + rx], 1u << ry);
                    else atomicOr(&sm.winO[i * Sp + rx], 1u << ry);
                }
            }
            }
            if (single_tool) {
                g.sync();
                for (int idx = lg; idx < N * S; idx += G) {
                    const int i = div_s(idx), r = idx - i * S;
                    sm.winF[i * Sp + r] = (r == R) ? (1u << R) : 0u;      // dec_grid_rl.py:233-234
                }
            }
        }
        g.sync();
    }

    // ---- fused memory update + observation rows ---------------------------------------------
    // Lane slot = (robot i, row a2 of the (2*rad+1)-row band around it), rad = max(ego, R).  The
    // <=2 interleaved {free_i, obst_i} words under the band row are loaded once; rows inside the
    // sensor range get the hits ORed in and written back; rows inside the egocentric window are
    // ORed into the observation bit stream.
    int cnt = 0;
    const int total = N * RW;
    const uint32_t wm_ow = low_mask(ow);
    const int ow2 = ow * ow;
    for (int idx = lg; idx < total; idx += G) {
        const int i = div_rw(idx), a2 = idx - i * RW;
        const uint32_t pi = sm.posF[i];
        const int xi = unpack_x(pi), yi = unpack_y(pi);
        const int x = xi - rad + a2;
        const bool inx = (unsigned)x < (unsigned)W;
        const int y0r = yi - rad, s0 = y0r >> 5, sh = y0r & 31;
        const bool in0 = inx && (unsigned)s0 < (unsigned)strips;
        const bool in1 = inx && (sh + RW > 32) && (unsigned)(s0 + 1) < (unsigned)strips;
        const uint32_t w0 = (uint32_t)(s0 * Wp + x);             // only dereferenced under in0 / in1
        const uint32_t fw = (eN + (uint32_t)i) * (uint32_t)PW + w0;
        const bool sens = do_sense && inx && abs(a2 - rad) <= R;
        const int ys0 = (yi - R) >> 5, ys1 = (yi + R) >> 5;
        const bool q0 = sens && in0, q1 = sens && in1;
        const bool t0 = q0 && s0 == ys0, t1 = q1 && s0 + 1 <= ys1;   // words the sensor rows touch

        uint2 lo = make_uint2(0, 0), hi = make_uint2(0, 0);
        uint32_t vlo = 0, vhi = 0;
        if (!fresh) {
            if (in0) lo = fo_g[fw];
            if (in1) hi = fo_g[fw + Wp];
            if (t0) vlo = vis_g[vis_off + w0];
            if (t1) vhi = vis_g[vis_off + w0 + Wp];
        }
        uint32_t f0 = 0, f1 = 0, o0 = 0, o1 = 0;
        if (!LIDAR) {
            // squaresensor.py:15-37 — the footprint is the truth map itself (sees through walls)
            const uint64_t m64 = (uint64_t)low_mask(S) << (sh + rad - R);
            const uint32_t m0 = (uint32_t)m64, m1 = (uint32_t)(m64 >> 32);
            if (q0) { o0 = ld_map<SP::SMT>(&wall_g[map_off + w0]) & m0; f0 = ld_map<SP::SMT>(&tpos_g[map_off + w0]) & m0; }
            if (q1) { o1 = ld_map<SP::SMT>(&wall_g[map_off + w0 + Wp]) & m1; f1 = ld_map<SP::SMT>(&tpos_g[map_off + w0 + Wp]) & m1; }
            if (single_tool) {                                     // dec_grid_rl.py:233-234
                const uint64_t c64 = (a2 == rad) ? (1ull << (sh + rad)) : 0ull;
                f0 = q0 ? (uint32_t)c64 : 0u;
                f1 = q1 ? (uint32_t)(c64 >> 32) : 0u;
            }
        } else if (sens) {
            const int r = a2 - rad + R;
            split_words(sm.winF[i * Sp + r] << (rad - R), y0r, f0, f1);
            split_words(sm.winO[i * Sp + r] << (rad - R), y0r, o0, o1);
            if (!in0) { f0 = 0; o0 = 0; }
            if (!in1) { f1 = 0; o1 = 0; }
        }
        // per-robot memory: free_i |= hits, obst_i |= hits  (dec_grid_rl.py:231-236)
        const uint2 nlo = make_uint2(lo.x | f0, lo.y | o0);
        const uint2 nhi = make_uint2(hi.x | f1, hi.y | o1);
        const int cn = __popc(f0 & ~lo.x) + __popc(f1 & ~hi.x);
        if ((f0 & ~lo.x) | (o0 & ~lo.y)) fo_g[fw] = nlo;
        if ((f1 & ~hi.x) | (o1 & ~hi.y)) fo_g[fw + Wp] = nhi;
        if (cn) atomicAdd(&sm.covS[i], cn);                         // shared-memory atomic
        // union "visited" plane: exactly one writer per word (dec_grid_rl.py:243-256)
        uint32_t comb0 = f0, comb1 = f1;
        bool own0 = t0, own1 = t1;
        const uint32_t nm = sens ? sm.nearM[i] : 0u;
        if (nm) {                                                   // rare: footprints overlap
            for (uint32_t mm = nm; mm; mm &= mm - 1) {
                const int j = __ffs(mm) - 1;
                const uint32_t pj = sm.posF[j];
                const int xj = unpack_x(pj), yj = unpack_y(pj);
                if (abs(x - xj) > R) continue;
                const int y0j = yj - R, s0j = y0j >> 5, s1j = (yj + R) >> 5;
#pragma unroll
                for (int k = 0; k < 2; ++k) {
                    const int sk = s0 + k;
                    if (!(k == 0 ? t0 : t1) || sk < s0j || sk > s1j) continue;
                    uint32_t add;
                    if (LIDAR) {
                        uint32_t jl, jh;
                        split_words(sm.winF[j * Sp + (x - xj + R)], y0j, jl, jh);
                        add = (sk == s0j) ? jl : jh;
                    } else if (single_tool) {
                        add = (x == xj && (yj >> 5) == sk) ? (1u << (yj & 31)) : 0u;
                    } else {
                        add = ld_map<SP::SMT>(&tpos_g[map_off + w0 + k * Wp]) & ymask_in_strip(yj, R, S, sk);
                    }
                    if (k == 0) { comb0 |= add; own0 = own0 && (j > i); }
                    else { comb1 |= add; own1 = own1 && (j > i); }
                }
            }
        }
        {
            const uint32_t nb0 = own0 ? (comb0 & ~vlo) : 0u, nb1 = own1 ? (comb1 & ~vhi) : 0u;
            cnt += __popc(nb0) + __popc(nb1);
            if (nb0) vis_g[vis_off + w0] = vlo | comb0;
            if (nb1) vis_g[vis_off + w0 + Wp] = vhi | comb1;
        }
        // egocentric observation rows (dec_grid_rl.py:334-347): layers 1 (free_i) and 2 (obst_i)
        const int a = a2 - (rad - ego);
        if ((unsigned)a < (unsigned)ow) {
            const uint32_t bF = (__funnelshift_r(nlo.x, nhi.x, sh) >> (rad - ego)) & wm_ow;
            const uint32_t bO = (__funnelshift_r(nlo.y, nhi.y, sh) >> (rad - ego)) & wm_ow;
            const int off = ((i * C + 1) * ow + a) * ow;
            stream_or(sm.bits, off, bF);
            stream_or(sm.bits, off + ow2, bO);
        }
    }
    // layer 0 shows every robot that has moved at least once (quirk A.3.4): one lane per
    // (viewer i, shown robot j) pair
    for (int idx = lg; idx < N * N; idx += G) {
        const int i = div_n(idx), j = idx - i * N;
        if (!((shown >> j) & 1u)) continue;
        const uint32_t pi = sm.posF[i], pj = sm.posF[j];
        const int a = unpack_x(pj) - (unpack_x(pi) - ego), b = unpack_y(pj) - (unpack_y(pi) - ego);
        if (a >= 0 && a < ow && b >= 0 && b < ow) stream_or(sm.bits, (i * C * ow + a) * ow + b, 1u);
    }
    g.sync();

    // ---- reward / termination bookkeeping ------------------------------------------------------
    cnt = g.sum(cnt);
    int covsum = 0;
    if (mode != MODE_OBSERVE) {
        int c = 0;
        if (lg < N) {
            const int dn = sm.covS[lg];
            c = cov_old + dn;
            if (fresh || dn != 0) hcov[lg] = c;                    // per-robot known-free count
        }
        covsum = g.sum(c);
    }
    bool again = false;
    if (mode == MODE_STEP) {
        int flag = 0;
        if (lg == 0) {
            double r = 0.0;
            int dn = 0, nr = 0;
            bool trunc = false;
            if (sentinel) {
                dn = 1;                                            // dec_grid_rl.py:105-107
                nr = 1;
            } else {
                for (int k = 0; k < ncoll; ++k) r -= p.cp;         // dec_grid_rl.py:185-204
                double obs_r = (double)cnt;
                if (!SP::FIXED && p.dist_term) obs_r = (double)p.dist_term[e] + (double)cnt;   // row D11
                r += obs_r;
                const int cs = cs0 + 1;                            // :154
                // percent_covered sums over robots (quirk A.3.3); min(thresh, 1) <= pct  (:156-157)
                bool full;
                if (th >= 1.0) full = covsum >= fc;                // exact: c/f >= 1  <=>  c >= f
                else full = th <= (double)covsum / (double)fc;
                if (full) { r += p.terminal; th += p.done_incr; dn = 1; head->done_thresh = th; }
                else if (cs == p.maxsteps) dn = 1;
                const double epr = epr0 + r;
                head->currstep = cs;
                head->flags = shown;
                head->ep_return = epr;
                trunc = p.truncate_steps > 0 && cs >= p.truncate_steps;
                sm.stats[4] += (double)ncoll;
                sm.stats[5] += (double)cnt;
                sm.stats[6] += (double)N;
                if (dn || trunc) {
                    sm.stats[0] += 1.0;
                    sm.stats[1] += epr;
                    sm.stats[2] += (double)cs;
                    sm.stats[3] += (double)covsum / (double)fc;
                    nr = 1;
                }
            }
            p.reward[e] = (float)r;
            if (p.reward64) p.reward64[e] = r;
            // bit 0: the reference's done(); bit 1: cut at truncate_steps (generate_episode's done = True,
            // Utils/utils.py:25-31) — visible to the caller even when the fused auto-reset consumes the episode end
            p.done[e] = (uint8_t)(dn | (trunc ? MCOV_DONE_TRUNCATED : 0));
            flag = nr && p.auto_reset;
            p.needs_reset[e] = (nr && !flag) ? 1 : 0;
        }
        again = g.shfl(flag, 0) != 0;
    } else if (fresh && lg == 0) {
        head->currstep = 0;                                        // :479
        head->map_idx = map_idx;
        head->flags = 0;                                           // _robot_pad starts empty (:487)
        head->episode = head->episode + 1;
        head->ep_return = 0.0;
        p.needs_reset[e] = 0;
    }
    if (again) return PASS_AGAIN;                                  // obs comes from the reset pass

    // ---- observation block: bit stream -> float32, coalesced 128-bit streaming stores ----------
    float* obs_e = p.obs + (size_t)e * p.obs_env_floats;
#if MCOV_PF_NEXT
    if (e_next >= 0) prefetch_next_env<SP>(p, nxt, e_next, g);
#endif
    if (SP::ST) {
        // hand the bit stream to the store warp: every lane's ORs into the slot are ordered before the
        // leader's store by the group barrier above; the store is a release at CTA scope
        if (lg == 0) {
            sm.bits[p.bits_words] = (p.debug & 1) ? 0xffffffffu : (uint32_t)e;
            st_release_smem(link.pub, (uint32_t)link.it + 1u);
        }
        return PASS_DONE;
    }
    if (p.debug & 1) return PASS_DONE;                             // MCOV_DEBUG=1: profiling aid, no obs store
    if (p.store_team > G) return PASS_STORE;                       // the block is stored by the group's team (kernel loop)
    if (p.vec_ok) {
        float4* dst = reinterpret_cast<float4*>(obs_e);
        const int nf4 = p.obs_env_floats >> 2;
#if MCOV_LUT8 == 1
        // one byte of the bit stream -> two float4 through a 256-entry table, two 128-bit stores per lane
        // (A/B on C5: 0.243 vs 0.196 ms — a lane's stores are 32 bytes apart, every instruction half-fills its sectors)
        {
            const int np = nf4 >> 1;
            float4* d2 = dst + 2 * lg;
            const int it2 = (np - lg + G - 1) / G;
MCOV_UNROLL(MCOV_ST_UNROLL)
            for (int t = 0; t < it2; ++t) {
                const int pq = lg + t * G;
                const uint32_t byte = (sm.bits[pq >> 2] >> ((pq & 3) * 8)) & 255u;
                const float4 v0 = lut8[2 * byte], v1 = lut8[2 * byte + 1];
                obs_store(d2 + 2 * t * G, v0);
                obs_store(d2 + 2 * t * G + 1, v1);
            }
            if ((nf4 & 1) && lg == 0) {
                const int q = nf4 - 1;
                obs_store(dst + q, lut[(sm.bits[q >> 3] >> ((q & 7) * 4)) & 15u]);
            }
            return PASS_DONE;
        }
#elif MCOV_LUT8 == 2
        // ... and one 256-bit store per lane (st.global.v8.f32, sm_100): a block that starts 16 bytes off a 32-byte
        // boundary (every other env at C2: 5808-byte blocks) peels its first float4, the byte is then cut from the
        // bit stream at a 4-bit phase
        {
            const int h = (int)((reinterpret_cast<uintptr_t>(dst) >> 4) & 1u);
            const int np = (nf4 - h) >> 1;
            if (lg == 0) {
                if (h) obs_store(dst, lut[sm.bits[0] & 15u]);
                if ((nf4 - h) & 1) {
                    const int q = nf4 - 1;
                    obs_store(dst + q, lut[(sm.bits[q >> 3] >> ((q & 7) * 4)) & 15u]);
                }
            }
            float4* d2 = dst + h + 2 * lg;
            const int it2 = (np - lg + G - 1) / G;
MCOV_UNROLL(MCOV_ST_UNROLL)
            for (int t = 0; t < it2; ++t) {
                const int pos = 4 * h + 8 * (lg + t * G);
                const uint32_t byte = __funnelshift_r(sm.bits[pos >> 5], sm.bits[(pos >> 5) + 1], pos & 31) & 255u;
                const float4 v0 = lut8[2 * byte], v1 = lut8[2 * byte + 1];
                asm volatile("st.global.cs.v8.f32 [%0], {%1, %2, %3, %4, %5, %6, %7, %8};"
                             :: "l"(d2 + 2 * t * G), "f"(v0.x), "f"(v0.y), "f"(v0.z), "f"(v0.w), "f"(v1.x), "f"(v1.y), "f"(v1.z), "f"(v1.w)
                             : "memory");
            }
            return PASS_DONE;
        }
#endif
        // float4 q = lg + G*t holds bits [4q, 4q+4): word q>>3, nibble q&7
        float4* d = dst + lg;
        const int iters = (nf4 - lg + G - 1) / G;
MCOV_UNROLL(MCOV_ST_UNROLL)
        for (int t = 0; t < iters; ++t) {
            const int q = lg + t * G;
            obs_store(d + t * G, lut[(sm.bits[q >> 3] >> ((q & 7) * 4)) & 15u]);
        }
    } else {
        for (int k = lg; k < p.obs_env_floats; k += G)
            __stcs(obs_e + k, ((sm.bits[k >> 5] >> (k & 31)) & 1u) ? 1.0f : 0.0f);
    }
    return PASS_DONE;
}

template <class SP>
__global__ void __launch_bounds__(SP::THREADS, SP::MINB)
dec_env_kernel(const DecParams p) {
    constexpr int G = SP::G;
    extern __shared__ __align__(16) unsigned char smem_raw[];
    __shared__ float4 lut[16];
#if MCOV_LUT8
    __shared__ float4 lut8[512];
    for (int t = threadIdx.x; t < 512; t += blockDim.x) {
        const uint32_t n = ((uint32_t)t >> 1), h = (n >> (4 * (t & 1))) & 15u;
        lut8[t] = make_float4((h & 1u) ? 1.0f : 0.0f, (h & 2u) ? 1.0f : 0.0f, (h & 4u) ? 1.0f : 0.0f, (h & 8u) ? 1.0f : 0.0f);
    }
#else
    const float4* lut8 = nullptr;
#endif
    __shared__ int last_flag;
    const Grp<G> g(threadIdx.x & 31);
    // Spec::ST: st_warps of the CTA's warps are store warps, the others are pass warps.  The store warps of the
    // CTAs resident on one SM must not share a warp scheduler (warp w issues on sub-partition w % 4; a store
    // warp carries a quarter of the SM's expansion work): CTA b takes warp (b / num_sms) % 4, i.e. the four CTAs
    // a persistent grid puts on an SM take four different sub-partitions.
    const int pass_threads = SP::ST ? (int)blockDim.x - 32 * p.st_warps : (int)blockDim.x;
    const int gpc = pass_threads / G;                     // lane groups (= envs in flight) per CTA
    const int wid = threadIdx.x >> 5;
    const int sw_first = SP::ST ? (p.st_rot_div > 0 ? (int)((blockIdx.x / (unsigned)p.st_rot_div) & 3u) : (pass_threads >> 5)) : 0;
    const bool store_warp = SP::ST && wid >= sw_first && wid < sw_first + p.st_warps;
    const int pass_tid = (SP::ST && wid >= sw_first + p.st_warps) ? (int)threadIdx.x - 32 * p.st_warps : (int)threadIdx.x;
    const int gi = store_warp ? 0 : pass_tid / G;
    GroupSmem sm = carve(smem_raw + (size_t)gi * p.grp_smem, p);
    sm.mapW = reinterpret_cast<const uint32_t*>(smem_raw + p.smt_map_off);
    sm.mapP = sm.mapW + (size_t)p.n_maps * p.g.PW;
    uint32_t* const bits0 = sm.bits;
    const uint32_t ctr0 = smem_u32(smem_raw + p.st_bar_off);   // uint32 pub[gpc] (blocks published), con[gpc] (consumed)

    if (threadIdx.x < 16) {
        const uint32_t t = threadIdx.x;
        lut[t] = make_float4((t & 1u) ? 1.0f : 0.0f, (t & 2u) ? 1.0f : 0.0f, (t & 4u) ? 1.0f : 0.0f,
                             (t & 8u) ? 1.0f : 0.0f);
    }
    if (!store_warp)
        for (int k = g.lg; k < MCOV_STATS_LEN; k += G) sm.stats[k] = 0.0;
    if (SP::ST)
        for (int k = threadIdx.x; k < gpc * 2; k += blockDim.x)
            reinterpret_cast<uint32_t*>(smem_raw + p.st_bar_off)[k] = 0u;
    uint64_t* const map_bar = reinterpret_cast<uint64_t*>(smem_raw + p.smt_bar_off);
    if (SP::SMT && threadIdx.x == 0) {
        mbar_init(map_bar, 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();
    if (SP::SMT) {
        // Map staging (TMA): the whole truth-map table — wall planes, then walkable planes, of every map — comes into
        // shared memory once per launch as bulk async copies of <= 32 kB signalled on one mbarrier; every env this
        // CTA steps afterwards reads its sensor footprint from there.
        if (threadIdx.x == 0) {
            const uint32_t plane_bytes = (uint32_t)p.n_maps * (uint32_t)p.g.PW * 4u;
            mbar_expect_tx(map_bar, 2u * plane_bytes);
            for (int t = 0; t < 2; ++t) {
                const unsigned char* src = reinterpret_cast<const unsigned char*>(t == 0 ? p.wall : p.pos);
                unsigned char* dst = smem_raw + p.smt_map_off + (size_t)t * plane_bytes;
                for (uint32_t o = 0; o < plane_bytes; o += 32768u)
                    bulk_load_g2s(dst + o, src + o, min(32768u, plane_bytes - o), map_bar);
            }
        }
        mbar_wait(map_bar, 0);
    }

    const int gtot = gridDim.x * gpc;
    const int mode = p.mode;
    if (!store_warp) {
    const int gg = blockIdx.x * gpc + gi;
    StoreLink link;
    link.pub = ctr0 + 4u * gi;
    link.con = ctr0 + 4u * (gpc + gi);
    link.it = 0;
    Prefetch nxt;
    nxt.pos = 0; nxt.act8 = 0; nxt.act64 = 0; nxt.map_idx = 0; nxt.shown = 0;
    if (mode != MODE_RESET && gg < p.E) nxt = load_prefetch(p, gg, g.lg);
    // Store teams (p.store_team = T > G lanes, a power of two <= 32): the T / G lane groups of a team run their
    // passes side by side, then all T lanes store the team's blocks one after the other — a block leaves in
    // 16 T-byte pieces instead of 16 G-byte ones (8 -> 16 lanes: tools/ubench/store_patterns.cu 4.65 -> 5.4 TB/s).
    // The loop is team-uniform: a group without an env (tail of the batch, masked reset) still meets its team.
    const int T = (!SP::ST && p.store_team > G) ? p.store_team : G;
    const int gpt = T / G, tpos = gi & (gpt - 1);                  // groups per team, this group's place in it
    const int lane = threadIdx.x & 31, tbase = lane & ~(T - 1);
    const unsigned tmask = T == 32 ? FULL : (((1u << (T & 31)) - 1u) << tbase);
    for (int e0 = gg - tpos; e0 < p.E; e0 += gtot) {
        const int e = e0 + tpos;
        int has_block = 0;
        if (e < p.E) {
        const Prefetch cur = nxt;
        if (mode != MODE_RESET && e + gtot < p.E) nxt = load_prefetch(p, e + gtot, g.lg);
        int m = mode;
        if (SP::ST) sm.bits = bits0 + (link.it & 1) * p.st_slot_words;
        bool act = true;
        if (m == MODE_RESET) {
            if (p.mask) act = p.mask[e] != 0;
            else if (p.mask_from_flags) act = p.needs_reset[e] != 0;
            if (!act && SP::ST) {                                  // group-uniform; the store warp counts slots: publish an empty one
                wait_slot_free(link);
                if (g.lg == 0) {
                    sm.bits[p.bits_words] = 0xffffffffu;
                    st_release_smem(link.pub, (uint32_t)link.it + 1u);
                }
                ++link.it;
            }
        }
        if (act && !SP::ST && (p.debug & 2)) {                     // MCOV_DEBUG=2: profiling aid, obs store only
            float4* d = reinterpret_cast<float4*>(p.obs + (size_t)e * p.obs_env_floats) + g.lg;
            const int iters = ((p.obs_env_floats >> 2) - g.lg + G - 1) / G;
#pragma unroll 4
            for (int t = 0; t < iters; ++t) {
                const int q = g.lg + t * G;
                obs_store(d + t * G, lut[(sm.bits[q >> 3] >> ((q & 7) * 4)) & 15u]);
            }
        } else if (act) {
            const int e_next = (mode == MODE_STEP && e + gtot < p.E) ? e + gtot : -1;
            int r;
            while ((r = env_pass<SP>(p, sm, lut, lut8, e, m, cur, g, nxt, e_next, link)) == PASS_AGAIN) m = MODE_RESET;   // fused auto-reset
            has_block = (r == PASS_STORE);
            ++link.it;
        }
        }
        if (T > G) {
            __syncwarp(tmask);                                     // the team's bit streams are complete
            const unsigned hb = __ballot_sync(tmask, has_block && g.lg == 0);
            const int tl = lane & (T - 1), nf4 = p.obs_env_floats >> 2;
            for (int b = 0; b < gpt; ++b) {
                if (!((hb >> (tbase + b * G)) & 1u)) continue;     // team-uniform
                const uint32_t* bb = reinterpret_cast<const uint32_t*>(smem_raw + (size_t)(gi - tpos + b) * p.grp_smem +
                                                                       MCOV_STATS_LEN * 8);
                float4* d = reinterpret_cast<float4*>(p.obs + (size_t)(e0 + b) * p.obs_env_floats);
MCOV_UNROLL(MCOV_ST_UNROLL)
                for (int q = tl; q < nf4; q += T)
                    obs_store(d + q, lut[(bb[q >> 3] >> ((q & 7) * 4)) & 15u]);
            }
            __syncwarp(tmask);                                     // ... and read out before the next pass clears them
        }
    }
    } else {
        // ---- store warp: full slot -> float32 staging buffer -> one TMA bulk store per env block ----
        // Store warp sw serves lane groups sw, sw + nsw, ...; its lane l tracks group l * nsw + sw: how many of
        // that group's blocks have been taken, how many it will publish.  One shared-memory load polls them all.
        const int lane = threadIdx.x & 31, nsw = p.st_warps, sw = wid - sw_first;
        const int mygrp = lane * nsw + sw;
        int taken = 0, mine = 0;
        if (mygrp < gpc) {
            const int gg = blockIdx.x * gpc + mygrp;
            if (gg < p.E) mine = (p.E - gg + gtot - 1) / gtot;
        }
        int left = (int)__reduce_add_sync(FULL, mine);
        const int nf4 = p.obs_env_floats >> 2;
        const uint32_t blk_bytes = (uint32_t)p.obs_env_floats * 4u;
        const int nbuf = p.st_nbuf;
        unsigned char* const stage0 = smem_raw + p.st_stage_off + (size_t)sw * nbuf * p.st_stage_pitch;
        const uint32_t my_pub = ctr0 + 4u * (mygrp < gpc ? mygrp : 0);
        int buf = 0;
        while (left > 0) {
            const bool ready = taken < mine && (int)ld_acquire_smem(my_pub) > taken;
            uint32_t rm = __ballot_sync(FULL, ready);
            if (!rm) { __nanosleep(32); continue; }
            for (; rm; rm &= rm - 1) {
                const int src = __ffs(rm) - 1, sgrp = src * nsw + sw;
                const int tk = __shfl_sync(FULL, taken, src);
                const uint32_t* bits = reinterpret_cast<const uint32_t*>(smem_raw + (size_t)sgrp * p.grp_smem + MCOV_STATS_LEN * 8) +
                                       (tk & 1) * p.st_slot_words;
                const uint32_t eid = bits[p.bits_words];
                if (eid != 0xffffffffu && !(p.debug & 4)) {          // MCOV_DEBUG=4: profiling aid, hand-over only
                    if (lane == 0) {                               // the staging buffer about to be refilled has been read out
                        if (nbuf == 4) asm volatile("cp.async.bulk.wait_group.read 3;" ::: "memory");
                        else if (nbuf == 3) asm volatile("cp.async.bulk.wait_group.read 2;" ::: "memory");
                        else asm volatile("cp.async.bulk.wait_group.read 1;" ::: "memory");
                    }
                    __syncwarp();
                    float4* stage = reinterpret_cast<float4*>(stage0 + (size_t)buf * p.st_stage_pitch);
                    // bits, table and staging buffer are all shared memory: the loads of a batch are issued before its
                    // stores by hand, or every store would order the next load behind it (possible aliasing)
                    const int sh4 = (lane & 7) * 4;
                    for (int q0 = lane; q0 < nf4; q0 += 32 * 6) {
                        uint32_t nib[6];
                        float4 v[6];
#pragma unroll
                        for (int j = 0; j < 6; ++j) {
                            const int q = q0 + 32 * j;
                            nib[j] = q < nf4 ? (bits[q >> 3] >> sh4) & 15u : 0u;
                        }
#pragma unroll
                        for (int j = 0; j < 6; ++j) v[j] = lut[nib[j]];
#pragma unroll
                        for (int j = 0; j < 6; ++j)
                            if (q0 + 32 * j < nf4) stage[q0 + 32 * j] = v[j];
                    }
                    fence_proxy_async_smem();                      // generic-proxy writes -> visible to the TMA engine
                    __syncwarp();
                    if (lane == 0) {
                        st_release_smem(ctr0 + 4u * (gpc + sgrp), (uint32_t)tk + 1u);   // slot is free again
                        if (!(p.debug & 8))                        // MCOV_DEBUG=8: profiling aid, expansion without the bulk store
                            bulk_store_s2g(p.obs + (size_t)eid * p.obs_env_floats, stage, blk_bytes);
                    }
                    buf = buf + 1 == nbuf ? 0 : buf + 1;
                } else {
                    __syncwarp();
                    if (lane == 0) st_release_smem(ctr0 + 4u * (gpc + sgrp), (uint32_t)tk + 1u);
                }
                if (lane == src) ++taken;
                --left;
            }
        }
        if (lane == 0) bulk_wait_all();                            // smem is read out and the blocks are written before the CTA ends
    }

    // ---- episode statistics: group sums -> CTA row -> last CTA folds all rows (single launch) --
    fold_episode_stats(p, smem_raw, gpc, (size_t)p.grp_smem, reinterpret_cast<double*>(smem_raw), &last_flag);
}

// ---- shareMaps (dec_grid_rl.py:423-447): free_i/obst_i <- OR over linked robots, then recount ---
__global__ void __launch_bounds__(256)
dec_share_kernel(int E, int N, int PW, unsigned char* hdr, int hdr_stride, int off_cov,
                 int off_adj, uint2* fo) {
    __shared__ int cnt[MCOV_MAX_ROBOTS];
    __shared__ uint32_t adj[MCOV_MAX_ROBOTS];
    for (int e = blockIdx.x; e < E; e += gridDim.x) {
        unsigned char* hrec = hdr + (size_t)e * hdr_stride;
        if (threadIdx.x < N) {
            adj[threadIdx.x] = reinterpret_cast<uint32_t*>(hrec + off_adj)[threadIdx.x] | (1u << threadIdx.x);
            cnt[threadIdx.x] = 0;
        }
        __syncthreads();
        bool any = false;
        for (int j = 0; j < N; ++j) any |= (adj[j] != (1u << j));
        if (any) {                                                 // CTA-uniform
            uint2* base = fo + (size_t)e * N * PW;
            for (int w = threadIdx.x; w < PW; w += blockDim.x) {
                uint2 v[MCOV_MAX_ROBOTS];
                for (int j = 0; j < N; ++j) v[j] = base[(size_t)j * PW + w];
                for (int r = 0; r < N; ++r) {
                    uint2 acc = make_uint2(0, 0);
                    const uint32_t a = adj[r];
                    for (int j = 0; j < N; ++j)
                        if ((a >> j) & 1u) { acc.x |= v[j].x; acc.y |= v[j].y; }
                    if (acc.x != v[r].x || acc.y != v[r].y) base[(size_t)r * PW + w] = acc;
                    const int c = __popc(acc.x);
                    if (c) atomicAdd(&cnt[r], c);                  // shared-memory atomic
                }
            }
            __syncthreads();
            if (threadIdx.x < N) reinterpret_cast<int32_t*>(hrec + off_cov)[threadIdx.x] = cnt[threadIdx.x];
        }
        __syncthreads();
    }
}

__global__ void stats_reduce_kernel(const double* partial, int rows, double* out) {
    const int k = threadIdx.x;
    if (k < MCOV_STATS_LEN) {
        double s = 0.0;
        for (int r = 0; r < rows; ++r) s += partial[(size_t)r * MCOV_STATS_LEN + k];
        out[k] = s;
    }
}

__global__ void init_headers_kernel(int E, unsigned char* hdr, int hdr_stride, double thresh,
                                    uint32_t flags0) {
    const int e = blockIdx.x * blockDim.x + threadIdx.x;
    if (e < E) {
        EnvHead* h = reinterpret_cast<EnvHead*>(hdr + (size_t)e * hdr_stride);
        h->done_thresh = thresh;
        h->ep_return = 0.0;
        h->currstep = 0;
        h->map_idx = 0;
        h->flags = flags0;
        h->episode = 0;
    }
}

__global__ void percent_covered_kernel(int E, int N, const unsigned char* hdr, int hdr_stride,
                                       int off_cov, const int32_t* free_counts, double* out) {
    const int e = blockIdx.x * blockDim.x + threadIdx.x;
    if (e < E) {
        const unsigned char* hrec = hdr + (size_t)e * hdr_stride;
        const EnvHead* h = reinterpret_cast<const EnvHead*>(hrec);
        const int32_t* cov = reinterpret_cast<const int32_t*>(hrec + off_cov);
        int s = 0;
        for (int j = 0; j < N; ++j) s += cov[j];
        out[e] = (double)s / (double)free_counts[h->map_idx];
    }
}

struct DecLaunchShape {
    int G, threads, grid;
    size_t grp_smem, smem;
    int bits_words;
    // store-warp variant (Spec::ST): 0 = off
    int st, slot_words, nbuf, bar_off, stage_off, stage_pitch, nsw;
    int team;                 // lanes that store a block together (>= G)
    int smt, map_off, map_bar_off;   // truth-map table resident in shared memory (Spec::SMT): byte offsets behind the group areas
};

inline int pick_group(const mcov_env* env) {
    const mcov_config& c = env->cfg;
    const int num_sms = env->num_sms;
    // lanes per env: at least the robots (poses live in lanes).  With many envs 8 lanes per env keep
    // the one-lane phases (moves, bookkeeping) shared by 4 envs per warp; with few envs (E * G below
    // the resident thread capacity) wider groups cut the per-env latency instead: a step of 4096
    // envs is one wave either way, so its duration is the latency of one env pass.
    int g = 8;
    if (c.num_robots > 16) g = 32;
    else if (c.num_robots > 8) g = 16;
    // the row loop of the band (robots x (2*rad+1) rows) should stay at <= 3 iterations: measured on B200
    // (profiles/r02_ab.jsonl), C5 (4 robots x 11 rows) 8 -> 16 lanes takes the step from 0.214 to 0.196 ms
    // (32 lanes: 0.226), C3 (8 robots x 21-row lidar bands) 8 -> 32 lanes from 0.160 to 0.111 ms
    const int rad = c.egoradius > c.sensor_range ? c.egoradius : c.sensor_range;
    const int rows = c.num_robots * (2 * rad + 1);
    while (g < 32 && rows > 3 * g) g *= 2;
    const long long capacity = (long long)num_sms * 4 * MAX_WARPS * 32;      // resident threads
    while (g < 32 && (long long)c.num_envs * (2 * g) <= capacity) g *= 2;
    // MCOV_DEC_GROUP=8|16|32 pins the width (tests run every width on the same traces; tuning)
    {
        const int v = env->pin_group;
        if ((v == 4 || v == 8 || v == 16 || v == 32) && v >= c.num_robots) g = v;
    }
    return g;
}

// Store-warp variant: parity-green, measured on B200 at C5 (profiles/r02_ab.jsonl) and SLOWER than the direct
// stores — 0.227-0.277 ms against 0.213 ms: the lane groups of a CTA publish in a convoy, one store warp expands
// about one block per microsecond, and the waiting groups spin (ncu: a third of all instructions in
// wait_slot_free).  It stays selectable (MCOV_DEC_STORE=1, read at mcov_create) as the TMA bulk-store arm of
// that A/B and is covered by the parity tests; the default is the direct path.
inline bool want_store_warp(const mcov_env* env, int G, bool vec_ok) {
    const mcov_config& c = env->cfg;
    const int ow = 2 * c.egoradius + 1;
    const long long blk = (long long)c.num_robots * env->obs_C * ow * ow * 4;
    if (!vec_ok || blk % 16 != 0 || blk > 12 * 1024 || G == 4) return false;
    return env->pin_store > 0;
}

inline DecLaunchShape dec_shape(const mcov_env* env, bool vec_ok) {
    const mcov_config& c = env->cfg;
    const int N = c.num_robots, S = 2 * c.sensor_range + 1, ow = 2 * c.egoradius + 1;
    DecLaunchShape s;
    s.G = pick_group(env);
    s.st = want_store_warp(env, s.G, vec_ok) ? 1 : 0;
    s.bits_words = (N * env->obs_C * ow * ow + 31) / 32 + 1;
    s.slot_words = s.st ? round_up(s.bits_words + 1, 2) : 0;       // + the env id word
    s.nbuf = 0; s.bar_off = 0; s.stage_off = 0; s.stage_pitch = 0; s.nsw = 1;
    int words = (s.st ? 2 * s.slot_words : s.bits_words) + 3 * N + (c.sensor_type == MCOV_SENSOR_LIDAR ? 3 * N * (S + 1) : 0);
    words = round_up(words, 2);
    // stagger the groups of one warp over the 32 banks: stride = multiple of 32 words + G
    int stride_words = MCOV_STATS_LEN * 2 + words;
    if (s.G < 32) stride_words = round_up(stride_words, 32) + s.G;
    s.grp_smem = (size_t)stride_words * 4;
    const int per_warp = 32 / s.G;
    int groups = per_warp, per_sm = 1;
    if (s.st) {
        // 7 pass warps + the store warp = 256 threads, the register budget of MCOV_DEC_MINB CTAs per SM
        const int blk = N * env->obs_C * ow * ow * 4;
        s.stage_pitch = round_up(blk, 128);
        s.nbuf = s.stage_pitch * 4 <= 24 * 1024 ? 4 : 2;
        if (env->pin_nbuf >= 2 && env->pin_nbuf <= 4) s.nbuf = env->pin_nbuf;
        s.nsw = env->pin_sw == 2 ? 2 : 1;
        int pw = MAX_WARPS - s.nsw;
        for (;; --pw) {
            groups = pw * per_warp;
            s.bar_off = (int)(s.grp_smem * groups);
            s.stage_off = round_up(s.bar_off + groups * 8, 128);
            s.smem = (size_t)s.stage_off + (size_t)s.nsw * s.nbuf * s.stage_pitch;
            if (s.smem <= 200 * 1024 || pw == 1) break;
        }
        s.threads = (pw + s.nsw) * 32;
        per_sm = (int)((220 * 1024) / (s.smem + 1024));
        const int by_regs = MCOV_DEC_MINB * MAX_WARPS / (pw + s.nsw);
        if (per_sm > by_regs) per_sm = by_regs;
        if (per_sm < 1) per_sm = 1;
    } else {
        // CTA size: the one that keeps the most lane groups resident per SM (shared memory is the limit
        // for lidar windows: 2 x 83 kB CTAs hold 64 groups where 5 x 41 kB CTAs hold 80)
        int best = 0;
        // lidar handles: 2-warp CTAs, measured on C3 (profiles/r02_ab.jsonl: 8 / 4 / 2 warps 0.1004 / 0.0984 / 0.0968 ms);
        // the square-sensor step prefers the largest CTA (C5: 8 / 4 / 2 warps 0.1959 / 0.1975 / 0.1984 ms)
        const int max_warps = (c.sensor_type == MCOV_SENSOR_LIDAR && env->pin_warps < 1) ? 2 : MAX_WARPS;
        for (int warps = max_warps; warps >= 1; warps >>= 1) {
            if (env->pin_warps >= 1 && warps != env->pin_warps && env->pin_warps <= MAX_WARPS) continue;   // MCOV_DEC_WARPS pins the CTA size
            const int gr = warps * per_warp;
            size_t sm_bytes = s.grp_smem * gr;
            if (sm_bytes < (size_t)warps * 32 * 8) sm_bytes = (size_t)warps * 32 * 8;
            if (sm_bytes > 200 * 1024) continue;
            int ps = (int)((220 * 1024) / (sm_bytes + 1024));
            const int by_regs = MCOV_DEC_MINB * MAX_WARPS / warps; // register budget of __launch_bounds__
            if (ps > by_regs) ps = by_regs;
            if (ps > 16) ps = 16;
            if (ps * gr > best) { best = ps * gr; groups = gr; per_sm = ps; }
        }
        // One wave (every env has a lane group resident at once): what counts is the busiest SM, so size the CTAs to
        // spread the batch's warps evenly — 4096 envs at 32 lanes are 4096 warps = 27.7 per SM: 512 CTAs of 8 warps put
        // 32 warps on some SMs and 24 on others, 586 CTAs of 7 warps put 28 on (almost) every one.
        if (env->pin_warps < 1 && env->pin_balance != 0) {
            const long long total_warps = ((long long)c.num_envs + per_warp - 1) / per_warp;
            const int cap_warps = per_sm * (groups / per_warp);           // resident warps per SM of the choice above
            const int wps = (int)((total_warps + env->num_sms - 1) / env->num_sms);
            if (wps <= cap_warps) {
                int best_w = groups / per_warp, best_waste = 1 << 30;
                for (int w = MAX_WARPS; w >= 1; --w) {
                    if (((w * 32) & 7) != 0) continue;
                    const int waste = (wps + w - 1) / w * w - wps;
                    if (waste < best_waste) { best_waste = waste; best_w = w; }
                }
                groups = best_w * per_warp;
                per_sm = (wps + best_w - 1) / best_w;
            }
        }
        s.threads = groups * s.G;
        s.smem = s.grp_smem * groups;
    }
    // Truth-map table in shared memory (Spec::SMT): square sensor, direct stores, a batch of several waves, and a table
    // that fits beside the group areas of one 1024-thread CTA per SM.  MCOV_DEC_SMT=0|1 pins it (read at mcov_create).
    s.smt = 0; s.map_off = 0; s.map_bar_off = 0;
    if (!s.st && c.sensor_type != MCOV_SENSOR_LIDAR && env->pin_warps < 1) {
        const size_t table = (size_t)env->n_maps * env->g.PW * 4 * 2;
        const int gr = 32 * per_warp;
        const size_t group_area = ((s.grp_smem * gr + 127) / 128) * 128;
        const bool fits = group_area + table + 128 <= 220 * 1024 && table % 16 == 0 &&
                          ((reinterpret_cast<uintptr_t>(env->wall) | reinterpret_cast<uintptr_t>(env->pos)) & 15u) == 0;
        bool on = false;                                           // default off until the A/B says otherwise
        if (env->pin_smt >= 0) on = env->pin_smt != 0;
        if (on && fits) {
            s.smt = 1;
            groups = gr;
            per_sm = 1;
            s.threads = 1024;
            s.map_off = (int)group_area;
            s.map_bar_off = (int)(group_area + table);
            s.smem = group_area + table + 128;
        }
    }
    // store teams: MCOV_DEC_TEAM=8|16|32 pins the lanes per block store (>= G; needs the vector path)
    s.team = s.G;
    if (!s.st && vec_ok && env->pin_team > s.G && env->pin_team <= 32 && (env->pin_team & (env->pin_team - 1)) == 0 &&
        groups % (env->pin_team / s.G) == 0) s.team = env->pin_team;
    // the statistics fold reuses the CTA's dynamic smem as blockDim doubles
    if (s.smem < (size_t)s.threads * 8) s.smem = (size_t)s.threads * 8;
    int grid = env->num_sms * per_sm;
    const int need = (c.num_envs + groups - 1) / groups;
    if (grid > need) grid = need;
    // MCOV_DEC_GRID=n caps the grid (tests: a few CTAs looping over many envs exercise the persistent loop
    // and the slot hand-over with small batches)
    if (env->pin_grid >= 1 && env->pin_grid < grid) grid = env->pin_grid;
    s.grid = grid;
    return s;
}


template <class SP>
cudaError_t launch_dec_t(mcov_env* env, const DecLaunchShape& s, const DecParams& p, cudaStream_t st) {
    // the opt-in is a per-device property of the kernel; a handle lives on one device and always launches the
    // same instantiation with the same shape, so it is set once per handle (not per process: a second device
    // in the same process needs its own call)
    const void* fn = reinterpret_cast<const void*>(&dec_env_kernel<SP>);
    if (s.smem > 48 * 1024 && (env->dec_smem_func != fn || env->dec_smem_configured < s.smem)) {
        cudaError_t err = cudaFuncSetAttribute(dec_env_kernel<SP>,
                                               cudaFuncAttributeMaxDynamicSharedMemorySize, (int)s.smem);
        if (err != cudaSuccess) return err;
        env->dec_smem_func = fn;
        env->dec_smem_configured = s.smem;
    }
    dec_env_kernel<SP><<<s.grid, s.threads, s.smem, st>>>(p);
    return cudaGetLastError();
}

}  // namespace

int dec_grid_size(const mcov_env* env) { return dec_shape(env, true).grid; }

// Largest grid a step launch of this handle can use, whichever kernel shape and store path it picks: the number of
// per-CTA statistics rows a launch may write (mcov_step_static_scratch sizes its check with it).
int dec_launch_grid_max(const mcov_env* env) {
    if (want_dec_cta(env)) return dec_cta_grid_size(env);
    const int a = dec_shape(env, true).grid, b = dec_shape(env, false).grid;
    return a > b ? a : b;
}

cudaError_t launch_dec(mcov_env* env, DecParams& p, cudaStream_t st) {
    if (want_dec_cta(env)) return launch_dec_cta(env, p, st);
    const DecLaunchShape s = dec_shape(env, p.vec_ok != 0 && p.obs != nullptr && (p.phases & PH_SENSE));   // a moves-only launch writes no block
    p.st_slot_words = s.slot_words;
    p.st_nbuf = s.nbuf;
    p.st_warps = s.nsw;
    p.st_rot_div = env->pin_rot == 0 ? 0 : env->num_sms;
    p.store_team = s.team;
    p.smt_map_off = s.map_off;
    p.smt_bar_off = s.map_bar_off;
    p.st_bar_off = s.bar_off;
    p.st_stage_off = s.stage_off;
    p.st_stage_pitch = s.stage_pitch;
    p.rad = p.ego > p.R ? p.ego : p.R;
    p.RW = 2 * p.rad + 1;
    p.bits_words = s.bits_words;
    p.grp_smem = (int)s.grp_smem;
    p.rw_magic = magic_for(p.RW);
    p.s_magic = magic_for(p.S);
    p.n_magic = magic_for(p.N);
    p.l_magic = magic_for(p.n_lasers > 0 ? p.n_lasers : 1);
    {
        const int nl = p.n_lasers > 0 ? p.n_lasers : 1;
        auto gcd = [](int a, int b) { while (b) { const int t = a % b; a = b; b = t; } return a; };
        int m = nl / 4 + 1;
        while (gcd(m, nl) != 1) ++m;
        p.l_mult = (uint32_t)(m % nl);
        if (nl == 1) p.l_mult = 0;
    }
    const bool lidar = p.sensor == MCOV_SENSOR_LIDAR;
    p.debug = env->debug;                                           // MCOV_DEBUG, read once at mcov_create
    if ((uint64_t)p.E * p.N * p.g.PW >= 0xffffffffull) return cudaErrorInvalidValue;   // 32-bit word offsets
    cudaError_t err;
    const bool plain = p.C == 3 && !p.single_tool && p.comm_radius == 0;
    // instantiations for the BASELINE configs (C2/C5, C3, C4), generic otherwise
    const bool c2 = plain && !lidar && p.N == 4 && p.R == 2 && p.ego == 5;
    const bool c3 = plain && lidar && p.N == 8 && p.R == 10 && p.ego == 5;
    const bool c4 = plain && !lidar && p.N == 16 && p.R == 2 && p.ego == 5;
    if (s.smt) {
        if (c2 && s.G == 16) err = launch_dec_t<Spec<16, false, 4, 2, 5, false, true>>(env, s, p, st);
        else if (s.G == 8) err = launch_dec_t<Spec<8, false, 0, 0, 0, false, true>>(env, s, p, st);
        else if (s.G == 16) err = launch_dec_t<Spec<16, false, 0, 0, 0, false, true>>(env, s, p, st);
        else err = launch_dec_t<Spec<32, false, 0, 0, 0, false, true>>(env, s, p, st);
    }
    else if (s.st) {
        if (c2 && s.G == 8) err = launch_dec_t<Spec<8, false, 4, 2, 5, true>>(env, s, p, st);
        else if (c2 && s.G == 32) err = launch_dec_t<Spec<32, false, 4, 2, 5, true>>(env, s, p, st);
        else if (c3 && s.G == 32) err = launch_dec_t<Spec<32, true, 8, 10, 5, true>>(env, s, p, st);
        else if (c4 && s.G == 16) err = launch_dec_t<Spec<16, false, 16, 2, 5, true>>(env, s, p, st);
        else if (s.G == 8) err = lidar ? launch_dec_t<Spec<8, true, 0, 0, 0, true>>(env, s, p, st) : launch_dec_t<Spec<8, false, 0, 0, 0, true>>(env, s, p, st);
        else if (s.G == 16) err = lidar ? launch_dec_t<Spec<16, true, 0, 0, 0, true>>(env, s, p, st) : launch_dec_t<Spec<16, false, 0, 0, 0, true>>(env, s, p, st);
        else err = lidar ? launch_dec_t<Spec<32, true, 0, 0, 0, true>>(env, s, p, st) : launch_dec_t<Spec<32, false, 0, 0, 0, true>>(env, s, p, st);
    }
    else if (c2 && s.G == 4) err = launch_dec_t<Spec<4, false, 4, 2, 5>>(env, s, p, st);
    else if (s.G == 4) err = lidar ? launch_dec_t<Spec<4, true, 0, 0, 0>>(env, s, p, st) : launch_dec_t<Spec<4, false, 0, 0, 0>>(env, s, p, st);
    else if (c2 && s.G == 8) err = launch_dec_t<Spec<8, false, 4, 2, 5>>(env, s, p, st);
    else if (c2 && s.G == 16) err = launch_dec_t<Spec<16, false, 4, 2, 5>>(env, s, p, st);
    else if (c2 && s.G == 32) err = launch_dec_t<Spec<32, false, 4, 2, 5>>(env, s, p, st);
    else if (c3 && s.G == 8) err = launch_dec_t<Spec<8, true, 8, 10, 5>>(env, s, p, st);
    else if (c3 && s.G == 32) err = launch_dec_t<Spec<32, true, 8, 10, 5>>(env, s, p, st);
    else if (c4 && s.G == 16) err = launch_dec_t<Spec<16, false, 16, 2, 5>>(env, s, p, st);
    else if (c4 && s.G == 32) err = launch_dec_t<Spec<32, false, 16, 2, 5>>(env, s, p, st);
    else if (s.G == 8) err = lidar ? launch_dec_t<Spec<8, true, 0, 0, 0>>(env, s, p, st) : launch_dec_t<Spec<8, false, 0, 0, 0>>(env, s, p, st);
    else if (s.G == 16) err = lidar ? launch_dec_t<Spec<16, true, 0, 0, 0>>(env, s, p, st) : launch_dec_t<Spec<16, false, 0, 0, 0>>(env, s, p, st);
    else err = lidar ? launch_dec_t<Spec<32, true, 0, 0, 0>>(env, s, p, st) : launch_dec_t<Spec<32, false, 0, 0, 0>>(env, s, p, st);
    env->launches++;
    return err;
}

cudaError_t launch_dec_share(mcov_env* env, cudaStream_t st) {
    const mcov_layout& l = env->lay;
    int grid = env->num_sms * 8;
    if (grid > env->cfg.num_envs) grid = env->cfg.num_envs;
    dec_share_kernel<<<grid, 256, 0, st>>>(env->cfg.num_envs, env->cfg.num_robots, l.plane_words,
                                          env->arena + l.off_hdr, l.hdr_stride, l.hdr_off_cov,
                                          l.hdr_off_adj, reinterpret_cast<uint2*>(env->arena + l.off_fo));
    env->launches++;
    return cudaGetLastError();
}

cudaError_t launch_stats_reduce(mcov_env* env, const double* partial, int rows, double* out,
                                cudaStream_t st) {
    stats_reduce_kernel<<<1, 32, 0, st>>>(partial, rows, out);
    env->launches++;
    return cudaGetLastError();
}

cudaError_t launch_init_headers(mcov_env* env, cudaStream_t st) {
    const int E = env->cfg.num_envs;
    const uint32_t flags0 = env->cfg.env_type == MCOV_ENV_SUPER ? 1u : 0u;
    init_headers_kernel<<<(E + 255) / 256, 256, 0, st>>>(E, env->arena + env->lay.off_hdr,
                                                         env->lay.hdr_stride, env->cfg.done_thresh, flags0);
    env->launches++;
    return cudaGetLastError();
}

cudaError_t launch_percent_covered(mcov_env* env, double* out, cudaStream_t st) {
    const int E = env->cfg.num_envs;
    const int ncov = env->cfg.env_type == MCOV_ENV_SUPER ? 1 : env->cfg.num_robots;
    percent_covered_kernel<<<(E + 255) / 256, 256, 0, st>>>(E, ncov, env->arena + env->lay.off_hdr,
                                                            env->lay.hdr_stride, env->lay.hdr_off_cov,
                                                            env->free_counts, out);
    env->launches++;
    return cudaGetLastError();
}

}  // namespace mcov
