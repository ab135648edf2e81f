// mcov_dec_cta.cu — DecGridRL reset/step/observe, CTA-per-env shape (small batches, large maps).
//
// Same semantics as dec_env_kernel (mcov_dec.cu; reference lines cited there: Environments/dec_grid_rl.py:91-258,
// 312-421, 449-552, Sensors/squaresensor.py:15-37, Sensors/lidar.py:16-65), different mapping: ONE CTA owns one env
// per loop iteration, warp w owns robot w, lane a owns row a of the robot's (2*rad+1)-row band.  A lane group of
// the other kernel walks N*(2*rad+1) window rows and N*C*ow*ow output floats with 8..32 lanes; here the same work
// is one row per lane and a few float4 per thread, so the duration of a small batch (256 envs x 16 robots on
// 1075x1075 maps: every env's CTA is resident at once) is a handful of dependent memory round trips instead of
// ~1300 warp instructions issued back to back by one warp.
//
// * Moves are resolved redundantly by every warp (lanes = robots, N rounds of shuffle + ballot): no hand-over.
// * Map-window staging (TMA): lane 0 of a warp issues cp.async.bulk copies of the <= 2 word runs under the band
//   in the robot's own {free, obst} plane, the union plane and the two truth planes into the warp's staging area
//   in shared memory, all signalled on one mbarrier; the band rows are then cut from shared memory.  For a
//   1075x1075 map that is a 32-row x 2-strip tile of a 146 kB plane per robot.  MCOV_DEC_TMA=0 selects plain
//   vector loads of the same words (A/B in profiles/r02_ab.jsonl).
// * Union reward: a word of the union plane touched by several robots has exactly one writer (lowest robot
//   index); the other robots' contributions are rebuilt from the truth plane (square sensor) or read from their
//   hit windows in shared memory (lidar), as in the lane-group kernel.  Counts are reduced with warp shuffles and
//   one shared-memory word per robot: no atomics to global memory.
// * The observation block of the env is expanded from a bit stream in shared memory by the whole CTA with
//   coalesced 128-bit streaming stores.
#include "mcov_internal.h"
#include "mcov_dec_shared.cuh"

#ifndef MCOV_LIDAR_RUNTAB
#define MCOV_LIDAR_RUNTAB 0                  // see mcov_dec.cu
#endif

namespace mcov {

namespace {

// a staged run covers the band rows plus one row either side (the tile is requested for the pre-move pose)
constexpr int FO_SLOTS = 36;                 // uint2 words per staged strip run of the {free, obst} plane
constexpr int U32_SLOTS = 40;                // uint32 words per staged strip run of a 1-bit plane
constexpr int STAGE_BYTES = 1536;            // per warp: 2 strips x (36 x 8 + 3 x 40 x 4)

struct CtaSmem {
    double* stats;                           // [MCOV_STATS_LEN] running sums of this CTA (thread 0)
    uint32_t* bits;                          // [bits_words] observation bit stream of the env
    uint32_t* posS;                          // [N] reset draw
    uint32_t* posF;                          // [N] poses after the moves
    int* cntW;                               // [N] newly visited union cells counted by robot w's warp
    int* covW;                               // [N] per-robot known-free counts after this step
    int* flags;                              // [0] re-seed this env (auto-reset), [1] map index of the reset draw
    uint32_t* winF;                          // lidar: [N*S] free hits, window-relative bits
    uint32_t* winO;                          //        [N*S] obstacle hits
    uint32_t* wallW;                         //        [N*S] truth wall window
    uint64_t* bars;                          // [N] one mbarrier per warp (TMA staging)
    unsigned char* stage;                    // [N][STAGE_BYTES]
};

__host__ __device__ inline size_t cta_carve(unsigned char* base, int N, int S, int bits_words, bool lidar, CtaSmem* s) {
    size_t off = 0;
    auto take = [&](size_t bytes, size_t align) { off = (off + align - 1) / align * align; const size_t o = off; off += bytes; return o; };
    const size_t o_stats = take(MCOV_STATS_LEN * 8, 16), o_bits = take((size_t)bits_words * 4, 16);
    const size_t o_pos = take((size_t)N * 4, 4), o_posf = take((size_t)N * 4, 4), o_cnt = take((size_t)N * 4, 4), o_cov = take((size_t)N * 4, 4);
    const size_t o_flags = take(16, 16);
    const size_t nwin = lidar ? (size_t)N * (S + 1) * 4 : 0;        // S window rows + the dummy row of the beam table walk
    const size_t o_wf = take(nwin, 4), o_wo = take(nwin, 4), o_ww = take(nwin, 4);
    const size_t o_bars = take((size_t)N * 8, 8), o_stage = take((size_t)N * STAGE_BYTES, 128);
    if (s) {
        s->stats = reinterpret_cast<double*>(base + o_stats);
        s->bits = reinterpret_cast<uint32_t*>(base + o_bits);
        s->posS = reinterpret_cast<uint32_t*>(base + o_pos);
        s->posF = reinterpret_cast<uint32_t*>(base + o_posf);
        s->cntW = reinterpret_cast<int*>(base + o_cnt);
        s->covW = reinterpret_cast<int*>(base + o_cov);
        s->flags = reinterpret_cast<int*>(base + o_flags);
        s->winF = reinterpret_cast<uint32_t*>(base + o_wf);
        s->winO = reinterpret_cast<uint32_t*>(base + o_wo);
        s->wallW = reinterpret_cast<uint32_t*>(base + o_ww);
        s->bars = reinterpret_cast<uint64_t*>(base + o_bars);
        s->stage = base + o_stage;
    }
    return off;
}

// -DMCOV_CTA_TICKS: thread 0 of every CTA records clock64() at the phase boundaries of its LAST pass into the
// statistics scratch rows (8 x int64 per CTA; tools/cta_ticks.py reads them) — a profiling build, never shipped.
#ifdef MCOV_CTA_TICKS
#define CTA_TICK(k) do { if (threadIdx.x == 0) reinterpret_cast<long long*>(p.stats_partial)[(size_t)blockIdx.x * 8 + (k)] = clock64(); } while (0)
#else
#define CTA_TICK(k) do { } while (0)
#endif

__device__ __forceinline__ int warp_sum(int v) { return (int)__reduce_add_sync(FULL, v); }

// Staged runs of one robot's band: row x of staged strip k sits at index x - xs (uint2 plane) / x - xs4 (1-bit planes).
constexpr int STRIP_BYTES = FO_SLOTS * 8 + 3 * U32_SLOTS * 4;
struct Staged {
    unsigned char* base;
    int xs, xs4;                             // first staged row of the uint2 / 1-bit runs
    int sb, xlo, xhi;                        // first staged strip, staged row interval
    __device__ __forceinline__ uint2* fo(int k) const { return reinterpret_cast<uint2*>(base + k * STRIP_BYTES); }
    __device__ __forceinline__ uint32_t* vis(int k) const { return reinterpret_cast<uint32_t*>(base + k * STRIP_BYTES + FO_SLOTS * 8); }
    __device__ __forceinline__ uint32_t* wall(int k) const { return vis(k) + U32_SLOTS; }
    __device__ __forceinline__ uint32_t* tpos(int k) const { return vis(k) + 2 * U32_SLOTS; }
};

__device__ __forceinline__ Staged staged_view(unsigned char* stage) {
    Staged v;
    v.base = stage;
    v.xs = v.xs4 = v.sb = v.xlo = 0;
    v.xhi = -1;
    return v;
}

// One pass over env e in `mode` by the whole CTA.  Returns PASS_AGAIN (CTA-uniform) when a finished env must be
// re-seeded in place by a second (MODE_RESET) pass (fused auto-reset).
template <bool LIDAR, bool TMA>
__device__ __forceinline__ int cta_pass(const DecParams& p, const CtaSmem& sm, const float4* lut, const int e,
                                        const int mode, uint32_t& tma_parity) {
    const int N = p.N, R = p.R, ego = p.ego, C = p.C;
    const bool single_tool = p.single_tool != 0;
    const int comm_radius = p.comm_radius, phases = p.phases;
    const int S = 2 * R + 1, Sp = S + 1, ow = 2 * ego + 1, rad = p.rad, RW = p.RW;
    const int W = p.g.W, L = p.g.L, Wp = p.g.Wp, strips = p.g.strips, PW = p.g.PW;
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;       // warp w <-> robot w

    unsigned char* hrec = p.hdr + (size_t)e * p.hdr_stride;
    EnvHead* head = reinterpret_cast<EnvHead*>(hrec);
    uint32_t* hpos = reinterpret_cast<uint32_t*>(hrec + p.off_pos);
    int32_t* hcov = reinterpret_cast<int32_t*>(hrec + p.off_cov);
    const uint32_t eN = (uint32_t)e * (uint32_t)N;
    const uint32_t vis_off = (uint32_t)e * (uint32_t)PW;
    uint2* const fo_g = p.fo;
    uint32_t* const vis_g = p.visited;

    __syncthreads();                             // the previous pass is done with the CTA's shared memory
    CTA_TICK(0);
    const bool fresh = (mode == MODE_RESET);
    int map_idx;
    uint32_t shown, myv;
    if (fresh) {
        if (threadIdx.x == 0) {
            int mi = 0;
            draw_reset(p, e, head->episode, sm.posS, &mi);
            sm.flags[1] = mi;
        }
        if (!p.skip_wipe) {
            // wipe this env's planes (dec_grid_rl.py:505-514) with 128-bit stores
            uint4* z = reinterpret_cast<uint4*>(fo_g + (size_t)eN * PW);
            const int n16 = N * (PW / 2);
            for (int k = threadIdx.x; k < n16; k += blockDim.x) z[k] = make_uint4(0, 0, 0, 0);
            uint4* zv = reinterpret_cast<uint4*>(vis_g + vis_off);
            for (int k = threadIdx.x; k < PW / 4; k += blockDim.x) zv[k] = make_uint4(0, 0, 0, 0);
        }
        __syncthreads();
        map_idx = sm.flags[1];
        myv = lane < N ? sm.posS[lane] : 0u;
        shown = 0;
    } else {
        map_idx = head->map_idx;
        shown = head->flags;
        myv = lane < N ? hpos[lane] : 0u;
    }
    const uint32_t map_off = (uint32_t)map_idx * (uint32_t)PW;
    const uint32_t* const wall_g = p.wall;
    const uint32_t* const tpos_g = p.pos;

    // ---- map-window staging (TMA): as soon as the pre-move pose is known, lane 0 requests the tile under the
    // band of robot w grown by one cell in every direction — wherever the move ends, the band is inside it — so
    // the copies run while the wall bits are fetched and the moves are resolved.
    Staged sv = staged_view(sm.stage + (size_t)w * STAGE_BYTES);
    uint32_t tma_bytes = 0;
    auto stage_tile = [&](int xc, int yc, int margin, bool want_fo, bool want_vis, bool want_truth) {
        sv.xlo = max(xc - p.rad - margin, 0);
        sv.xhi = min(xc + p.rad + margin, W - 1);
        sv.xs = sv.xlo & ~1;
        sv.xs4 = sv.xlo & ~3;
        const int yb = yc - p.rad - margin;
        sv.sb = yb >> 5;
        const bool k1 = (yb & 31) + p.RW + 2 * margin > 32;
        const bool h0 = (unsigned)sv.sb < (unsigned)strips, h1 = k1 && (unsigned)(sv.sb + 1) < (unsigned)strips;
        const uint32_t nfo = (uint32_t)(((sv.xhi + 2) & ~1) - sv.xs) * 8u, n32 = (uint32_t)(((sv.xhi + 4) & ~3) - sv.xs4) * 4u;
        const uint32_t bytes = (uint32_t)((h0 ? 1 : 0) + (h1 ? 1 : 0)) *
                               ((want_fo ? nfo : 0u) + (want_vis ? n32 : 0u) + (want_truth ? 2u * n32 : 0u));
        if (lane == 0 && bytes) {
            uint64_t* bar = sm.bars + w;
            mbar_expect_tx(bar, bytes);
#pragma unroll
            for (int k = 0; k < 2; ++k) {
                if (!(k == 0 ? h0 : h1)) continue;
                const uint32_t row = (uint32_t)((sv.sb + k) * Wp);
                if (want_fo) bulk_load_g2s(sv.fo(k), fo_g + (size_t)(eN + w) * PW + row + sv.xs, nfo, bar);
                if (want_vis) bulk_load_g2s(sv.vis(k), vis_g + vis_off + row + sv.xs4, n32, bar);
                if (want_truth) {
                    bulk_load_g2s(sv.wall(k), wall_g + map_off + row + sv.xs4, n32, bar);
                    bulk_load_g2s(sv.tpos(k), tpos_g + map_off + row + sv.xs4, n32, bar);
                }
            }
        }
        return bytes;                            // warp-uniform
    };
    const bool st_fo = !fresh, st_vis = !fresh && mode == MODE_STEP, st_truth = !LIDAR && (fresh || mode == MODE_STEP);
    if (TMA && !(mode == MODE_STEP && !(phases & PH_SENSE))) {      // a moves-only launch reads no window
        const uint32_t pw_ = __shfl_sync(FULL, myv, w);
        tma_bytes = stage_tile(unpack_x(pw_), unpack_y(pw_), (mode == MODE_STEP && (phases & PH_MOVE)) ? 1 : 0,
                               st_fo, st_vis, st_truth);
    }

    // early loads for the bookkeeping at the end of the pass
    int cs0 = 0, fc = 1;
    double th = 0.0, epr0 = 0.0;
    if (mode == MODE_STEP && threadIdx.x == 0) {
        cs0 = head->currstep;
        th = head->done_thresh;
        epr0 = head->ep_return;
        fc = __ldg(p.free_counts + map_idx);
    }
    int cov_old = 0;
    if (mode == MODE_STEP && lane == 0) cov_old = hcov[w];

    int px = unpack_x(myv), py = unpack_y(myv);
    int ncoll = 0;
    bool sentinel = false;
    if (mode == MODE_STEP) {
        // ---- action decode (dec_grid_rl.py:110-117); every warp resolves all moves, lanes = robots ----
        int u;
        if (p.act_fmt == MCOV_ACT_I64_BASE4) {
            const long long a64 = reinterpret_cast<const long long*>(p.actions)[e];
            sentinel = (a64 == -1);
            u = (int)((a64 >> (2 * lane)) & 3);
        } else {
            u = lane < N ? (int)(reinterpret_cast<const uint8_t*>(p.actions)[(size_t)e * N + lane] & 3u) : 0;
        }
        if (!sentinel && (phases & PH_MOVE)) {
            // sequential, collision-checked moves (dec_grid_rl.py:128-145,171-204); 0:x+1 1:y+1 2:x-1 3:y-1
            const int tx = px + (u == 0) - (u == 2);
            const int ty = py + (u == 1) - (u == 3);
            int blocked0 = 1;
            if (lane < N && tx >= 0 && tx < W && ty >= 0 && ty < L) blocked0 = get_bit(wall_g + map_off, Wp, tx, ty);
            const uint32_t tv = pack_xy(tx, ty);
            uint32_t pv = myv;
#ifdef MCOV_CTA_TICKS
            if (__ballot_sync(FULL, tx > 1000000)) return PASS_DONE;   // waits for poses and actions
            CTA_TICK(1);
            if (__ballot_sync(FULL, blocked0 > 1)) return PASS_DONE;    // waits for the wall bits
            CTA_TICK(2);
#endif
            // Fast path (N <= 16): when no target cell equals any robot's cell or another robot's target, the order of
            // the moves does not matter — every robot that is not blocked by a wall moves.  One MATCH over the 2N
            // cells (lanes 0..15 current cells, lanes 16..31 targets; blocked / unused lanes hold unique sentinels).
            bool sequential = true;
            if (N <= 16) {
                uint32_t val = 0xffff0000u | (uint32_t)lane;
                if (lane < N) val = myv;
                const uint32_t tsrc = blocked0 ? (0xfffe0000u | (uint32_t)lane) : tv;
                const uint32_t tk = __shfl_sync(FULL, tsrc, lane & 15);
                if (lane >= 16 && lane - 16 < N) val = tk;
                const uint32_t same = __match_any_sync(FULL, val);
                sequential = __any_sync(FULL, (same & (same - 1u)) != 0u) != 0;
            }
            if (!sequential) {
                const uint32_t movers = __ballot_sync(FULL, lane < N && !blocked0);
                if ((movers >> lane) & 1u) pv = tv;
                shown |= movers;
                ncoll = N - __popc(movers);
            } else {
            for (int k = 0; k < N; ++k) {
                const uint32_t tvk = __shfl_sync(FULL, tv, k);
                const int blk = __shfl_sync(FULL, blocked0, k);
                const uint32_t b = __ballot_sync(FULL, (lane < N) && pv == tvk);
                if (!blk && b == 0) {
                    if (lane == k) pv = tvk;
                    shown |= 1u << k;
                } else {
                    ++ncoll;
                }
            }
            }
            px = unpack_x(pv);
            py = unpack_y(pv);
        } else if (!sentinel) {
            const uint2 ms = p.move_scratch[e];                    // resolved by an earlier PH_MOVE launch
            ncoll = (int)ms.x;
            shown = ms.y;
        }
    }
    const uint32_t newv = pack_xy(px, py);
    CTA_TICK(3);
    for (int k = threadIdx.x; k < p.bits_words; k += blockDim.x) sm.bits[k] = 0;
    if (w == 0 && lane < N) sm.posF[lane] = newv;

    // robot of this warp
    const int i = w;
    const int xi = __shfl_sync(FULL, px, i), yi = __shfl_sync(FULL, py, i);

    // ---- communication graph (dec_grid_rl.py:374-391): row i by warp i ---------------------------------
    if (mode != MODE_OBSERVE && (comm_radius > 0 || fresh)) {
        const uint32_t b = __ballot_sync(FULL, (lane < N) && max(abs(px - xi), abs(py - yi)) <= comm_radius);
        if (lane == 0) reinterpret_cast<uint32_t*>(hrec + p.off_adj)[i] = comm_radius > 0 ? b : (1u << i);
    }
    const bool persist = (w == 0 && lane < N && mode != MODE_OBSERVE && (fresh || newv != myv));
    if (mode == MODE_STEP && !(phases & PH_SENSE)) {
        __syncthreads();                         // every warp has read the old poses
        if (persist) hpos[lane] = newv;
        if (threadIdx.x == 0) p.move_scratch[e] = make_uint2((uint32_t)ncoll, shown);
        return PASS_DONE;
    }
    const bool do_sense = fresh || (mode == MODE_STEP && !sentinel);

    // ---- band geometry of robot i -------------------------------------------------------------------------
    const int y0r = yi - rad, s0 = y0r >> 5, sh = y0r & 31;
    const bool two = sh + RW > 32;
    const bool sin0 = (unsigned)s0 < (unsigned)strips, sin1 = two && (unsigned)(s0 + 1) < (unsigned)strips;
    const int xa = max(xi - rad, 0), xb = min(xi + rad, W - 1);

    __syncthreads();                             // bit stream zeroed before any row is ORed in; old poses read by every warp
    if (persist) hpos[lane] = newv;

    // which robots' sensor rows can share a bitmap word with robot i's?
    uint32_t nm = 0;
    if (do_sense) {
        const bool nr = (lane < N) && (lane != i) && abs(xi - px) <= 2 * R &&
                        max((yi - R) >> 5, (py - R) >> 5) <= min((yi + R) >> 5, (py + R) >> 5);
        nm = __ballot_sync(FULL, nr);
    }
    if (LIDAR && do_sense) {
        // lidar.py:16-65 — stage the truth wall window of robot i, then march its beams (lanes = beams)
        uint32_t* ww = sm.wallW + i * Sp;
        uint32_t* wf = sm.winF + i * Sp;
        uint32_t* wo = sm.winO + i * Sp;
        if (lane == 0) ww[S] = 0xffffffffu;                          // dummy row: all wall
        if (lane < S) {
            const int x = xi - R + lane, y0 = yi - R;
            const uint32_t wrow = (x >= 0 && x < W) ? read_window(wall_g + map_off, Wp, strips, x, y0, S) : 0xffffffffu;
            ww[lane] = wrow;
            const bool own_wall = (wrow >> R) & 1u;                // see dec_env_kernel: a probe may stand on an obstacle
            wf[lane] = (p.lidar_cells && lane == R && !own_wall) ? (1u << R) : 0u;
            wo[lane] = (p.lidar_cells && lane == R && own_wall) ? (1u << R) : 0u;
        }
        __syncwarp();
        if (p.lidar_cells) {
            if (!((ww[R] >> R) & 1u)) {
                for (int l = lane; l < p.n_lasers; l += 32) {
#if MCOV_LIDAR_RUNTAB
                    walk_beam_runs(p.lidar_runs, l, ww, wf, wo);
#else
                    walk_beam_cells(p.lidar_cells, l, ww, wf, wo);
#endif
                }
            }
        } else {
            const double rng = (double)R;
            for (int l = lane; l < p.n_lasers; l += 32) {
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
                    walk = inside && inwin && !((ww[rx] >> ry) & 1u);
                    if (!(walk && dist < rng)) break;
                    atomicOr(&wf[rx], 1u << ry);
                    cx += xinc;
                    cy += yinc;
                    dist += dinc;
                }
                if (rx >= 0 && rx < S && ry >= 0 && ry < S) {
                    if (walk) atomicOr(&wf[rx], 1u << ry);
                    else atomicOr(&wo[rx], 1u << ry);
                }
            }
        }
        if (single_tool) {
            __syncwarp();
            if (lane < S) wf[lane] = (lane == R) ? (1u << R) : 0u;  // dec_grid_rl.py:233-234
        }
        __syncthreads();                         // other robots' hit windows are read for the union plane
    }

    // ---- fused memory update + observation row: lane a2 owns row a2 of robot i's band ----------------------
    if (TMA) {
        if (tma_bytes) {
            mbar_wait(sm.bars + w, tma_parity);  // the staged runs have landed
            tma_parity ^= 1u;
        }
        // the band after the move lies inside the requested tile, except when a 31-row band (radius 15) grown by
        // its margin would need a third strip: request the exact tile then
        const bool covered = xa >= sv.xlo && xb <= sv.xhi && s0 >= sv.sb && s0 + (two ? 1 : 0) <= sv.sb + 1;
        if (!covered) {
            tma_bytes = stage_tile(xi, yi, 0, st_fo, st_vis, st_truth);
            if (tma_bytes) {
                mbar_wait(sm.bars + w, tma_parity);
                tma_parity ^= 1u;
            }
        }
    }
    const int kb = TMA ? s0 - sv.sb : 0;          // staged strip slot of the band's first strip
    CTA_TICK(4);
    int cnt = 0, cn = 0;
    const uint32_t wm_ow = low_mask(ow);
    const int ow2 = ow * ow;
    if (lane < RW) {
        const int a2 = lane;
        const int x = xi - rad + a2;
        const bool inx = (unsigned)x < (unsigned)W;
        const bool in0 = inx && sin0, in1 = inx && sin1;
        const uint32_t w0 = (uint32_t)(s0 * Wp + x);             // only dereferenced under in0 / in1
        const uint32_t fw = (eN + (uint32_t)i) * (uint32_t)PW + w0;
        const bool sens = do_sense && inx && abs(a2 - rad) <= R;
        const int ys0 = (yi - R) >> 5, ys1 = (yi + R) >> 5;
        const bool q0 = sens && in0, q1 = sens && in1;
        const bool t0 = q0 && s0 == ys0, t1 = q1 && s0 + 1 <= ys1;   // words the sensor rows touch

        uint2 lo = make_uint2(0, 0), hi = make_uint2(0, 0);
        uint32_t vlo = 0, vhi = 0;
        if (!fresh) {
            if (TMA) {
                if (in0) lo = sv.fo(kb)[x - sv.xs];
                if (in1) hi = sv.fo(1)[x - sv.xs];               // in1 implies kb == 0 (covered)
                if (t0) vlo = sv.vis(kb)[x - sv.xs4];
                if (t1) vhi = sv.vis(1)[x - sv.xs4];
            } else {
                if (in0) lo = fo_g[fw];
                if (in1) hi = fo_g[fw + Wp];
                if (t0) vlo = vis_g[vis_off + w0];
                if (t1) vhi = vis_g[vis_off + w0 + Wp];
            }
        }
        uint32_t f0 = 0, f1 = 0, o0 = 0, o1 = 0, tp0 = 0, tp1 = 0;
        if (!LIDAR) {
            // squaresensor.py:15-37 — the footprint is the truth map itself (sees through walls)
            const uint64_t m64 = (uint64_t)low_mask(S) << (sh + rad - R);
            const uint32_t m0 = (uint32_t)m64, m1 = (uint32_t)(m64 >> 32);
            if (q0) {
                tp0 = TMA ? sv.tpos(kb)[x - sv.xs4] : __ldg(&tpos_g[map_off + w0]);
                o0 = (TMA ? sv.wall(kb)[x - sv.xs4] : __ldg(&wall_g[map_off + w0])) & m0;
                f0 = tp0 & m0;
            }
            if (q1) {
                tp1 = TMA ? sv.tpos(1)[x - sv.xs4] : __ldg(&tpos_g[map_off + w0 + Wp]);
                o1 = (TMA ? sv.wall(1)[x - sv.xs4] : __ldg(&wall_g[map_off + w0 + Wp])) & m1;
                f1 = tp1 & m1;
            }
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
        cn = __popc(f0 & ~lo.x) + __popc(f1 & ~hi.x);
        if ((f0 & ~lo.x) | (o0 & ~lo.y)) fo_g[fw] = nlo;
        if ((f1 & ~hi.x) | (o1 & ~hi.y)) fo_g[fw + Wp] = nhi;
        // union "visited" plane: exactly one writer per word (dec_grid_rl.py:243-256)
        uint32_t comb0 = f0, comb1 = f1;
        bool own0 = t0, own1 = t1;
        if (sens && nm) {                                          // rare: footprints overlap
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
                        add = (k == 0 ? tp0 : tp1) & ymask_in_strip(yj, R, S, sk);
                    }
                    if (k == 0) { comb0 |= add; own0 = own0 && (j > i); }
                    else { comb1 |= add; own1 = own1 && (j > i); }
                }
            }
        }
        {
            const uint32_t nb0 = own0 ? (comb0 & ~vlo) : 0u, nb1 = own1 ? (comb1 & ~vhi) : 0u;
            cnt = __popc(nb0) + __popc(nb1);
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
    // layer 0 shows every robot that has moved at least once (quirk A.3.4): lane j = shown robot j, viewer i
    if (lane < N && ((shown >> lane) & 1u)) {
        const int a = px - (xi - ego), b = py - (yi - ego);
        if (a >= 0 && a < ow && b >= 0 && b < ow) stream_or(sm.bits, (i * C * ow + a) * ow + b, 1u);
    }
    cnt = warp_sum(cnt);
    cn = warp_sum(cn);
    if (lane == 0) {
        sm.cntW[i] = cnt;
        int c = 0;
        if (mode != MODE_OBSERVE) {
            c = cov_old + cn;
            if (fresh || cn != 0) hcov[i] = c;                    // per-robot known-free count
        }
        sm.covW[i] = c;
    }
    __syncthreads();
    CTA_TICK(5);

    // ---- reward / termination bookkeeping (warp 0 folds the per-robot counts, thread 0 decides) ------------
    if (w == 0) {
        const int cnt_all = warp_sum(lane < N ? sm.cntW[lane] : 0);
        const int covsum = warp_sum(lane < N ? sm.covW[lane] : 0);
        if (lane == 0) {
            int flag = 0;
            if (mode == MODE_STEP) {
                double r = 0.0;
                int dn = 0, nr = 0;
                bool trunc = false;
                if (sentinel) {
                    dn = 1;                                        // dec_grid_rl.py:105-107
                    nr = 1;
                } else {
                    for (int k = 0; k < ncoll; ++k) r -= p.cp;     // dec_grid_rl.py:185-204
                    double obs_r = (double)cnt_all;
                    if (p.dist_term) obs_r = (double)p.dist_term[e] + (double)cnt_all;   // row D11
                    r += obs_r;
                    const int cs = cs0 + 1;                        // :154
                    bool full;                                     // percent_covered sums over robots (quirk A.3.3)
                    if (th >= 1.0) full = covsum >= fc;            // exact: c/f >= 1  <=>  c >= f
                    else full = th <= (double)covsum / (double)fc;
                    if (full) { r += p.terminal; th += p.done_incr; dn = 1; head->done_thresh = th; }
                    else if (cs == p.maxsteps) dn = 1;
                    const double epr = epr0 + r;
                    head->currstep = cs;
                    head->flags = shown;
                    head->ep_return = epr;
                    trunc = p.truncate_steps > 0 && cs >= p.truncate_steps;
                    sm.stats[4] += (double)ncoll;
                    sm.stats[5] += (double)cnt_all;
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
                p.done[e] = (uint8_t)(dn | (trunc ? MCOV_DONE_TRUNCATED : 0));
                flag = nr && p.auto_reset;
                p.needs_reset[e] = (nr && !flag) ? 1 : 0;
            } else if (fresh) {
                head->currstep = 0;                                // :479
                head->map_idx = map_idx;
                head->flags = 0;                                   // _robot_pad starts empty (:487)
                head->episode = head->episode + 1;
                head->ep_return = 0.0;
                p.needs_reset[e] = 0;
            }
            sm.flags[0] = flag;
        }
    }
    CTA_TICK(6);

    // ---- observation block: bit stream -> float32, coalesced 128-bit streaming stores by the whole CTA -----
    // (not behind the bookkeeping: the other warps store while thread 0 decides; when the env turns out to be
    // finished and is re-seeded in place, the reset pass overwrites the block after the barrier below)
    float* obs_e = p.obs + (size_t)e * p.obs_env_floats;
    if (p.debug & 1) {                                             // MCOV_DEBUG=1: profiling aid, no obs store
    } else if (p.vec_ok) {
        float4* dst = reinterpret_cast<float4*>(obs_e);
        const int nf4 = p.obs_env_floats >> 2;
        for (int q = threadIdx.x; q < nf4; q += blockDim.x)
            obs_store(dst + q, lut[(sm.bits[q >> 3] >> ((q & 7) * 4)) & 15u]);
    } else {
        for (int k = threadIdx.x; k < p.obs_env_floats; k += blockDim.x)
            __stcs(obs_e + k, ((sm.bits[k >> 5] >> (k & 31)) & 1u) ? 1.0f : 0.0f);
    }
    CTA_TICK(7);
    __syncthreads();
    return sm.flags[0] ? PASS_AGAIN : PASS_DONE;
}

template <bool LIDAR, bool TMA>
__global__ void __launch_bounds__(1024)
dec_cta_kernel(const DecParams p) {
    extern __shared__ __align__(128) unsigned char smem_raw[];
    __shared__ float4 lut[16];
    __shared__ int last_flag;
    CtaSmem sm;
    cta_carve(smem_raw, p.N, p.S, p.bits_words, LIDAR, &sm);
    if (threadIdx.x < 16) {
        const uint32_t t = threadIdx.x;
        lut[t] = make_float4((t & 1u) ? 1.0f : 0.0f, (t & 2u) ? 1.0f : 0.0f, (t & 4u) ? 1.0f : 0.0f,
                             (t & 8u) ? 1.0f : 0.0f);
    }
    if (threadIdx.x < MCOV_STATS_LEN) sm.stats[threadIdx.x] = 0.0;
    if (TMA && (threadIdx.x & 31) == 0) mbar_init(sm.bars + (threadIdx.x >> 5), 1);
    if (TMA) asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    __syncthreads();

    uint32_t tma_parity = 0;
    for (int e = blockIdx.x; e < p.E; e += gridDim.x) {
        int m = p.mode;
        if (m == MODE_RESET) {
            int act = 1;
            if (p.mask) act = p.mask[e];
            else if (p.mask_from_flags) act = p.needs_reset[e];
            if (!act) continue;                                    // CTA-uniform
        }
        while (cta_pass<LIDAR, TMA>(p, sm, lut, e, m, tma_parity) == PASS_AGAIN) m = MODE_RESET;   // fused auto-reset
    }
    fold_episode_stats(p, reinterpret_cast<const unsigned char*>(sm.stats), 1, 0, reinterpret_cast<double*>(smem_raw),
                       &last_flag);
}

struct CtaShape {
    int threads, grid;
    size_t smem;
    int bits_words;
};

inline CtaShape cta_shape(const mcov_env* env) {
    const mcov_config& c = env->cfg;
    const int N = c.num_robots, S = 2 * c.sensor_range + 1, ow = 2 * c.egoradius + 1;
    CtaShape s;
    s.bits_words = (N * env->obs_C * ow * ow + 31) / 32 + 1;
    s.threads = N * 32;
    s.smem = cta_carve(nullptr, N, S, s.bits_words, c.sensor_type == MCOV_SENSOR_LIDAR, nullptr);
    if (s.smem < (size_t)s.threads * 8) s.smem = (size_t)s.threads * 8;   // the statistics fold reuses it as blockDim doubles
    int per_sm = 2048 / s.threads;                                  // 64 registers per thread: 2048 threads per SM
    const int by_smem = (int)((220 * 1024) / (s.smem + 1024));
    if (per_sm > by_smem) per_sm = by_smem;
    if (per_sm < 1) per_sm = 1;
    int grid = env->num_sms * per_sm;
    if (grid > c.num_envs) grid = c.num_envs;
    if (env->pin_grid >= 1 && env->pin_grid < grid) grid = env->pin_grid;
    s.grid = grid;
    return s;
}

template <bool LIDAR, bool TMA>
cudaError_t launch_cta_t(mcov_env* env, const CtaShape& s, const DecParams& p, cudaStream_t st) {
    const void* fn = reinterpret_cast<const void*>(&dec_cta_kernel<LIDAR, TMA>);
    if (s.smem > 48 * 1024 && (env->dec_smem_func != fn || env->dec_smem_configured < s.smem)) {
        cudaError_t err = cudaFuncSetAttribute(dec_cta_kernel<LIDAR, TMA>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                               (int)s.smem);
        if (err != cudaSuccess) return err;
        env->dec_smem_func = fn;
        env->dec_smem_configured = s.smem;
    }
    dec_cta_kernel<LIDAR, TMA><<<s.grid, s.threads, s.smem, st>>>(p);
    return cudaGetLastError();
}

}  // namespace

// The CTA-per-env shape pays when every env's CTA is resident at once (E x N warps fit the machine): then the
// step lasts as long as one pass.  Beyond that the lane-group kernel's throughput wins.  MCOV_DEC_SHAPE=0|1
// (read at mcov_create) pins the choice: tests run every golden trace through both.
bool want_dec_cta(const mcov_env* env) {
    if (env->pin_shape >= 0) return env->pin_shape == 1;
    const mcov_config& c = env->cfg;
    return c.num_robots >= 2 && (long long)c.num_envs * c.num_robots * 32 <= (long long)env->num_sms * 2048;
}

int dec_cta_grid_size(const mcov_env* env) { return cta_shape(env).grid; }

cudaError_t launch_dec_cta(mcov_env* env, DecParams& p, cudaStream_t st) {
    const CtaShape s = cta_shape(env);
    p.st_slot_words = 0;
    p.store_team = 0;
    p.smt_map_off = p.smt_bar_off = 0;
    p.st_nbuf = p.st_warps = p.st_rot_div = p.st_bar_off = p.st_stage_off = p.st_stage_pitch = 0;
    p.rad = p.ego > p.R ? p.ego : p.R;
    p.RW = 2 * p.rad + 1;
    p.bits_words = s.bits_words;
    p.grp_smem = 0;
    p.rw_magic = p.s_magic = p.n_magic = p.l_magic = 0;
    p.debug = env->debug;
    if ((uint64_t)p.E * p.N * p.g.PW >= 0xffffffffull) return cudaErrorInvalidValue;   // 32-bit word offsets
    const bool lidar = p.sensor == MCOV_SENSOR_LIDAR;
    // bulk copies need 16-byte aligned sources: the arena planes are (mcov_state_layout), a caller's map table may not be
    const bool tma = env->pin_tma != 0 && ((reinterpret_cast<uintptr_t>(p.wall) | reinterpret_cast<uintptr_t>(p.pos)) & 15u) == 0;
    cudaError_t err;
    if (lidar) err = tma ? launch_cta_t<true, true>(env, s, p, st) : launch_cta_t<true, false>(env, s, p, st);
    else err = tma ? launch_cta_t<false, true>(env, s, p, st) : launch_cta_t<false, false>(env, s, p, st);
    env->launches++;
    return err;
}

}  // namespace mcov
