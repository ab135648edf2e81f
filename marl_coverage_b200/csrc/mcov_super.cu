// mcov_super.cu — SuperGridRL reset/step/get_state as ONE fused sm_100a kernel per call.
//
// Reference semantics restated here (file:line in ExistentialRobotics/MARL-Coverage):
//   SuperGridRL.step          Environments/super_grid_rl.py:74-225
//   motion_penalty            Environments/super_grid_rl.py:227-243   (first-step quirk)
//   isInBounds / isOccupied   Environments/super_grid_rl.py:245-279
//   get_distance_map          Environments/super_grid_rl.py:281-303   (cv2 DIST_L1, restated exactly)
//   get_state / get_pos_image Environments/super_grid_rl.py:305-341
//   reset                     Environments/super_grid_rl.py:343-399
//   done / percent_covered    Environments/super_grid_rl.py:401-421
//
// Mapping: one CTA owns one env per loop iteration (persistent CTA-stride loop).  Warp 0 resolves
// the N sequential collision-checked moves (poses live in lanes, shuffle + ballot), then stamps the
// (2r+1)^2 footprints robot by robot (lanes = footprint rows) into the bit-packed `covered` /
// `seen_obst` planes and counts the newly covered cells with __popc; the float64 reward is
// accumulated in the reference's raster order so it is bit-identical even with a fractional
// free_penalty.  The whole CTA then rebuilds the L1 distance channel (closed-form bit pass along y,
// min-plus sweep along x on a uint16 work buffer in shared memory) and streams the (N+3) x W x L
// float32 state — 280 kB per env at 4 agents / 100x100, the HBM-write-bound part — with 128-bit
// streaming stores straight into the torch buffer.  Finished envs are re-seeded by the same CTA in
// the same launch (fused auto-reset); episode statistics are folded by the last CTA.
#include "mcov_internal.h"
#include "mcov_dt.cuh"

namespace mcov {

namespace {

constexpr unsigned FULL = 0xffffffffu;
#ifndef MCOV_SUPER_MINB
#define MCOV_SUPER_MINB 3                     // resident CTAs per SM the register budget is set for
#endif
constexpr int ST = 512;                       // threads per CTA
constexpr int MAXS = 2 * MCOV_MAX_RADIUS + 1;

struct SuperSmem {
    double stats[MCOV_STATS_LEN];             // running sums of this CTA (thread 0 only)
    double rew[MCOV_MAX_ROBOTS];              // per control-slot rewards (super_grid_rl.py:104)
    float4 lut[16];                           // 4 bits -> float4
    uint2 rows[MCOV_MAX_ROBOTS * MAXS];       // per (robot, footprint row): .x new cells worth 1, .y re-covered
    uint32_t pos[MCOV_MAX_ROBOTS];            // poses after the moves
    int any_zero, dmax, covered, again, skip, last;
    int work;                                 // next unclaimed writer item of the current env
};

// dynamic shared memory: [dt scratch ints][distance LUT floats: W + L + 2][distance work buffer]
struct SuperDyn {
    int* red;
    float* dlut;
    uint16_t* dbuf;
};

// Distance transform of the covered plane: zero pixels are the NOT-covered in-map cells
// (obstacles included) — super_grid_rl.py:292-295.
// Also fills the value table dlut[d] = 1 - d / max(d) (float32, super_grid_rl.py:297-303) so that the
// state writer and the reward lookup never divide.
__device__ void super_transform(const uint32_t* cov, const Geometry& g, const SuperDyn& dyn, int pitch,
                                SuperSmem& sm, const Team team) {
    const int Wp = g.Wp, strips = g.strips, lastbits = g.L & 31;
    const DtResult r = l1_transform(
        [&](int ss, int x) -> uint32_t {
            const uint32_t valid = (ss == strips - 1 && lastbits) ? low_mask(lastbits) : FULL;
            return ~cov[ss * Wp + x] & valid;
        }, g, dyn.dbuf, pitch, dyn.red, team);
    if (team.rank == 0) { sm.any_zero = r.any_zero; sm.dmax = r.dmax; }
    if (r.any_zero)
        for (int k = team.rank; k <= r.dmax; k += team.size) dyn.dlut[k] = dist_value(k, 1, r.dmax);
    team.sync();
}

// np.sum of a float64 vector the way numpy adds it (pairwise_sum with 8 accumulators for n >= 8;
// verified against numpy 2.3 for every n <= 32) — super_grid_rl.py:214.
__device__ double numpy_sum(const double* a, int n) {
    if (n < 8) {
        double r = 0.0;
        for (int i = 0; i < n; ++i) r = __dadd_rn(r, a[i]);
        return r;
    }
    double r[8];
    for (int k = 0; k < 8; ++k) r[k] = a[k];
    int i = 8;
    for (; i < n - (n % 8); i += 8)
        for (int k = 0; k < 8; ++k) r[k] = __dadd_rn(r[k], a[i + k]);
    double res = __dadd_rn(__dadd_rn(__dadd_rn(r[0], r[1]), __dadd_rn(r[2], r[3])),
                           __dadd_rn(__dadd_rn(r[4], r[5]), __dadd_rn(r[6], r[7])));
    for (; i < n; ++i) res = __dadd_rn(res, a[i]);
    return res;
}

// Per-env hand-over from the dynamics warp to the state writers (double buffered).
struct EnvSlot {
    uint32_t pos[MCOV_MAX_ROBOTS];            // poses after the moves
    int covered;                              // covered-cell count after the step (0: nothing to transform)
};

// ---- state tensor (super_grid_rl.py:305-341) ------------------------------------------------------
// [pos_0 .. pos_{N-1} | one merged plane with use_scanning] [observed obstacles] [free = NOT covered]
// [distance channel], each W x L float32, y fastest.  Written by threads [first, first + nthreads)
// of the CTA; no barrier inside: the thread that zero-fills the float4 holding a robot's cell also
// patches the 1.0 into it (same thread, same address: program order).
__device__ void write_state(const SuperParams& p, const SuperSmem& sm, const EnvSlot& slot, const SuperDyn& dy,
                            int e, int first, int nthreads) {
    const int rank = (int)threadIdx.x - first;
    if (rank < 0 || rank >= nthreads) return;
    const uint16_t* dbuf = dy.dbuf;
    const float* dlut = dy.dlut;
    const int N = p.N, W = p.g.W, L = p.g.L, Wp = p.g.Wp;
    const int WL = W * L, pitch = p.dt_pitch;
    const int npos = p.use_scanning ? 1 : N;
    const uint32_t* cov = p.covered + (size_t)e * p.g.PW;
    const uint32_t* seen = p.seen_obst + (size_t)e * p.g.PW;
    // nothing covered (or no division, :299): channel is all 1; no zero pixel at all: all 0 (cv2 FLT_MAX)
    const bool flat = slot.covered == 0 || !sm.any_zero || sm.dmax <= 0;
    const float flat_v = (slot.covered == 0 || sm.any_zero) ? 1.0f : 0.0f;
    float* obs_e = p.obs + (size_t)e * p.C * WL;
    if (p.vec_ok) {
        const int Lq = L >> 2, qp = WL >> 2;
        float4* dst = reinterpret_cast<float4*>(obs_e);
        const float4 z4 = make_float4(0.f, 0.f, 0.f, 0.f);
        const int nz = npos * qp;
        for (int q = rank; q < nz; q += nthreads) __stcs(dst + q, z4);
        for (int j = 0; j < N; ++j) {
            const uint32_t pv = slot.pos[j];
            const int cell = unpack_x(pv) * L + unpack_y(pv);
            const int q = (p.use_scanning ? 0 : j) * qp + (cell >> 2);
            if (q % nthreads == rank) obs_e[(size_t)(p.use_scanning ? 0 : j) * WL + cell] = 1.0f;
        }
        float4* d_obst = dst + (size_t)npos * qp;
        float4* d_free = d_obst + qp;
        float4* d_dist = d_free + qp;
        int x = rank / Lq, yq = rank - x * Lq;
        const int dx = nthreads / Lq, dyq = nthreads - dx * Lq;
        for (int q = rank; q < qp; q += nthreads) {
            const int w = (yq >> 3) * Wp + x, nib = (yq & 7) * 4;
            const uint32_t so = seen[w], cv = cov[w];
            __stcs(d_obst + q, sm.lut[(so >> nib) & 15u]);
            __stcs(d_free + q, sm.lut[(~cv >> nib) & 15u]);
            float4 dv;
            if (flat) {
                dv = make_float4(flat_v, flat_v, flat_v, flat_v);
            } else {
                const uint32_t* dp = reinterpret_cast<const uint32_t*>(dbuf + (size_t)x * pitch + 4 * yq);
                const uint32_t a = dp[0], b = dp[1];
                dv.x = dlut[a & 0xffffu];
                dv.y = dlut[a >> 16];
                dv.z = dlut[b & 0xffffu];
                dv.w = dlut[b >> 16];
            }
            __stcs(d_dist + q, dv);
            x += dx;
            yq += dyq;
            if (yq >= Lq) { yq -= Lq; ++x; }
        }
    } else {
        // generic path (L not a multiple of 4 or unaligned buffer): one float per thread
        const int total = p.C * WL;
        for (int k = rank; k < total; k += nthreads) {
            const int c = k / WL, rem = k - c * WL;
            const int x = rem / L, y = rem - x * L;
            float v;
            if (c < npos) {
                v = 0.0f;
                if (p.use_scanning) {
                    for (int j = 0; j < N; ++j) v = (slot.pos[j] == pack_xy(x, y)) ? 1.0f : v;
                } else {
                    v = (slot.pos[c] == pack_xy(x, y)) ? 1.0f : 0.0f;
                }
            } else {
                const int w = (y >> 5) * Wp + x, b = y & 31;
                if (c == npos) v = (float)((seen[w] >> b) & 1u);
                else if (c == npos + 1) v = (float)((~cov[w] >> b) & 1u);
                else v = flat ? flat_v : dlut[dbuf[(size_t)x * pitch + y]];
            }
            __stcs(obs_e + k, v);
        }
    }
}

// ---- overlapped schedule (vectorised path) ---------------------------------------------------------
// The six planes that do not depend on the distance transform (positions, observed obstacles, free)
// are cut into warp items of IT float4; warps claim items from a shared counter, so the warps that
// run the transform simply join the writers when they are done.
#ifndef MCOV_SUPER_IT
#define MCOV_SUPER_IT 160
#endif
constexpr int IT = MCOV_SUPER_IT;                            // float4 per warp item (IT / 32 per lane)

__device__ void write_fixed_planes(const SuperParams& p, SuperSmem& sm, const EnvSlot& slot, int e, int lane) {
    const int N = p.N, W = p.g.W, L = p.g.L, Wp = p.g.Wp;
    const int WL = W * L, Lq = L >> 2, qp = WL >> 2;
    const int npos = p.use_scanning ? 1 : N, nz = npos * qp;
    const int n1 = (nz + IT - 1) / IT, n2 = (qp + IT - 1) / IT;
    const uint32_t* cov = p.covered + (size_t)e * p.g.PW;
    const uint32_t* seen = p.seen_obst + (size_t)e * p.g.PW;
    float* obs_e = p.obs + (size_t)e * p.C * WL;
    float4* dst = reinterpret_cast<float4*>(obs_e);
    float4* d_obst = dst + (size_t)nz;
    float4* d_free = d_obst + qp;
    const float4 z4 = make_float4(0.f, 0.f, 0.f, 0.f);
    const int dx = 32 / Lq, dyq = 32 - dx * Lq;
    while (true) {
        int it = 0;
        if (lane == 0) it = atomicAdd(&sm.work, 1);
        it = __shfl_sync(FULL, it, 0);
        if (it >= n1 + n2) break;
        if (it < n1) {
            const int base = it * IT;
#pragma unroll
            for (int t = 0; t < IT / 32; ++t) {
                const int q = base + lane + 32 * t;
                if (q < nz) __stcs(dst + q, z4);
            }
            // lane j patches robot j's 1.0 into the float4 another lane just zero-filled: __syncwarp orders
            // the two stores (one cell per robot, N <= 32)
            __syncwarp();
            if (lane < N) {
                const uint32_t pv = slot.pos[lane];
                const int cell = unpack_x(pv) * L + unpack_y(pv);
                const int plane = p.use_scanning ? 0 : lane;
                const int q = plane * qp + (cell >> 2);
                if (q >= base && q < base + IT) obs_e[(size_t)plane * WL + cell] = 1.0f;
            }
        } else {
            const int base = (it - n1) * IT;
            int q = base + lane;
            int x = q / Lq, yq = q - x * Lq;
#pragma unroll
            for (int t = 0; t < IT / 32; ++t) {
                if (q < qp) {
                    const int w = (yq >> 3) * Wp + x, nib = (yq & 7) * 4;
                    const uint32_t so = seen[w], cv = cov[w];
                    __stcs(d_obst + q, sm.lut[(so >> nib) & 15u]);
                    __stcs(d_free + q, sm.lut[(~cv >> nib) & 15u]);
                }
                q += 32;
                x += dx;
                yq += dyq;
                if (yq >= Lq) { yq -= Lq; ++x; }
            }
        }
    }
}

// The distance plane, statically split over threads [first, first + nthreads).
__device__ void write_dist_plane(const SuperParams& p, const SuperSmem& sm, const EnvSlot& slot,
                                 const SuperDyn& dy, int e, int first, int nthreads) {
    const int rank = (int)threadIdx.x - first;
    if (rank < 0 || rank >= nthreads) return;
    const int W = p.g.W, L = p.g.L, WL = W * L, Lq = L >> 2, qp = WL >> 2, pitch = p.dt_pitch;
    const int npos = p.use_scanning ? 1 : p.N;
    const bool flat = slot.covered == 0 || !sm.any_zero || sm.dmax <= 0;
    const float flat_v = (slot.covered == 0 || sm.any_zero) ? 1.0f : 0.0f;
    float4* d_dist = reinterpret_cast<float4*>(p.obs + (size_t)e * p.C * WL) + (size_t)(npos + 2) * qp;
    if (flat) {
        const float4 fv = make_float4(flat_v, flat_v, flat_v, flat_v);
        for (int q = rank; q < qp; q += nthreads) __stcs(d_dist + q, fv);
        return;
    }
    const uint16_t* dbuf = dy.dbuf;
    const float* dlut = dy.dlut;
    int x = rank / Lq, yq = rank - x * Lq;
    const int dx = nthreads / Lq, dyq = nthreads - dx * Lq;
    for (int q = rank; q < qp; q += nthreads) {
        const uint32_t* dp = reinterpret_cast<const uint32_t*>(dbuf + (size_t)x * pitch + 4 * yq);
        const uint32_t a = dp[0], b = dp[1];
        __stcs(d_dist + q, make_float4(dlut[a & 0xffffu], dlut[a >> 16], dlut[b & 0xffffu], dlut[b >> 16]));
        x += dx;
        yq += dyq;
        if (yq >= Lq) { yq -= Lq; ++x; }
    }
}

// Episode start for env e, by warp 0 alone (super_grid_rl.py:343-399): new map, new poses, nothing
// covered, nothing sensed.  Does not touch a_prev / done_thresh (they persist).
__device__ void super_reset_env(const SuperParams& p, EnvSlot& slot, int e, int lane) {
    const int N = p.N, PW = p.g.PW;
    unsigned char* hrec = p.hdr + (size_t)e * p.hdr_stride;
    EnvHead* head = reinterpret_cast<EnvHead*>(hrec);
    uint32_t* hpos = reinterpret_cast<uint32_t*>(hrec + p.off_pos);
    int32_t* hcov = reinterpret_cast<int32_t*>(hrec + p.off_cov);
    if (lane == 0) {
        int mi = 0;
        draw_starts(p.wall, p.g, N, p.n_maps, p.map_idx_in, p.start_xy_in, p.seed,
                    (uint32_t)(p.env_id_base + e), e, head->episode, slot.pos, &mi);
        head->currstep = 0;
        head->map_idx = mi;
        head->episode = head->episode + 1;
        head->ep_return = 0.0;
        hcov[0] = 0;
        for (int j = 0; j < N; ++j) hpos[j] = slot.pos[j];
        p.needs_reset[e] = 0;
        slot.covered = 0;
    }
    uint4* zc = reinterpret_cast<uint4*>(p.covered + (size_t)e * PW);
    uint4* zs = reinterpret_cast<uint4*>(p.seen_obst + (size_t)e * PW);
    for (int k = lane; k < PW / 4; k += 32) {
        zc[k] = make_uint4(0, 0, 0, 0);
        zs[k] = make_uint4(0, 0, 0, 0);
    }
    __syncwarp();
}

// Dynamics of env e by warp 0 alone: action decode, scan order, collision-checked moves, footprints,
// reward, termination, statistics and — when the episode ended and auto_reset is on — the re-seed of
// the env in place.  Leaves poses + covered count in `slot` for the state writers.
__device__ void super_dynamics(const SuperParams& p, SuperSmem& sm, EnvSlot& slot, const SuperDyn& dy,
                               const int e, const int mode, const int lane) {
    uint16_t* const dbuf = dy.dbuf;
    const int N = p.N, R = p.R, S = 2 * R + 1;
    const int W = p.g.W, L = p.g.L, Wp = p.g.Wp, strips = p.g.strips, PW = p.g.PW;
    unsigned char* hrec = p.hdr + (size_t)e * p.hdr_stride;
    EnvHead* head = reinterpret_cast<EnvHead*>(hrec);
    uint32_t* hpos = reinterpret_cast<uint32_t*>(hrec + p.off_pos);
    int32_t* hcov = reinterpret_cast<int32_t*>(hrec + p.off_cov);
    uint32_t* cov = p.covered + (size_t)e * PW;
    uint32_t* seen = p.seen_obst + (size_t)e * PW;
    if (mode == MODE_RESET) {
        super_reset_env(p, slot, e, lane);
        return;
    }
    bool sentinel = false;
    if (mode == MODE_STEP && p.act_fmt == MCOV_ACT_I64_BASE4)
        sentinel = (reinterpret_cast<const long long*>(p.actions)[e] == -1);
    const bool dyn = (mode == MODE_STEP) && !sentinel;
    int again = 0;
    {
    const int map_idx = head->map_idx;
    const uint32_t* wall = p.wall + (size_t)map_idx * PW;
    const uint32_t* tpos = p.pos + (size_t)map_idx * PW;
    const uint32_t myv = lane < N ? hpos[lane] : 0xffffffffu;
    uint32_t pv = myv;
    if (dyn) {
        // ---- action decode (:92-100): digit c of the joint action drives control slot c ----
        int u = 0;
        if (p.act_fmt == MCOV_ACT_I64_BASE4)
            u = (int)((reinterpret_cast<const long long*>(p.actions)[e] >> (2 * lane)) & 3);
        else if (lane < N)
            u = (int)(reinterpret_cast<const uint8_t*>(p.actions)[(size_t)e * N + lane] & 3u);
        // ---- scan order (:107-111,125-129): control c goes to the c-th robot in raster order of
        // the start-of-step poses (key x + y*W, ties by index)
        int rank = lane, zc = lane;
        if (p.use_scanning) {
            const int key = unpack_x(myv) + unpack_y(myv) * W;
            rank = 0;
            for (int j = 0; j < N; ++j) {
                const int kj = __shfl_sync(FULL, key, j);
                rank += (kj < key) || (kj == key && j < lane);
            }
            for (int j = 0; j < N; ++j)
                if (__shfl_sync(FULL, rank, j) == lane) zc = j;
        }
        // ---- sequential, collision-checked moves (:121-174).  Super table: 0:x-1 1:x+1 2:y+1 3:y-1.
        // A robot moves only on its own turn, so control c's target is known up front.
        const uint32_t pz = __shfl_sync(FULL, myv, zc & 31);
        const int tx = unpack_x(pz) + (u == 1) - (u == 0);
        const int ty = unpack_y(pz) + (u == 2) - (u == 3);
        int blocked0 = 1;
        if (lane < N && tx >= 0 && tx < W && ty >= 0 && ty < L) blocked0 = get_bit(wall, Wp, tx, ty);
        const uint32_t tv = pack_xy(tx, ty);
        int moved = 0, ncoll = 0;
        for (int c = 0; c < N; ++c) {
            const uint32_t tvc = __shfl_sync(FULL, tv, c);
            const int blk = __shfl_sync(FULL, blocked0, c);
            const int z = __shfl_sync(FULL, zc, c);
            const uint32_t b = __ballot_sync(FULL, lane < N && pv == tvc);
            if (!blk && b == 0u) {
                if (lane == z) pv = tvc;
                if (lane == c) moved = 1;
            } else {
                ++ncoll;
            }
        }
        // per-slot reward so far: -collision_penalty, or the start-of-step distance value (:138-139)
        double rc = 0.0;
        if (lane < N) {
            if (!moved) rc = 0.0 - p.cp;
            else if (p.dist_reward)
                rc = 0.0 + (double)dist_value((int)dbuf[(size_t)tx * p.dt_pitch + ty], sm.any_zero, sm.dmax);
        }
        // robot i accumulates onto the slot that drove it (r2c, :129)
        const double racc0_lo = rc;
        const int hi32 = __shfl_sync(FULL, __double2hiint(racc0_lo), rank & 31);
        const int lo32 = __shfl_sync(FULL, __double2loint(racc0_lo), rank & 31);
        double racc = __hiloint2double(hi32, lo32);

        // ---- footprints (:176-201), robots in index order, lanes = footprint rows -------------
        const int lastbits = L & 31;
        int cnt_new = 0;
        for (int i = 0; i < N; ++i) {
            const uint32_t pi = __shfl_sync(FULL, pv, i);
            const int xi = unpack_x(pi), yi = unpack_y(pi);
            const int x = xi - R + lane;
            uint32_t rel_new = 0, rel_al = 0;
            if (lane < S && x >= 0 && x < W) {
                const int y0 = yi - R, s0 = y0 >> 5, sh = y0 & 31;
                const uint64_t m64 = (uint64_t)low_mask(S) << sh;
#pragma unroll
                for (int k = 0; k < 2; ++k) {
                    const int s = s0 + k;
                    uint32_t m = k ? (uint32_t)(m64 >> 32) : (uint32_t)m64;
                    if (s == strips - 1 && lastbits) m &= low_mask(lastbits);
                    if (!m || s < 0 || s >= strips) continue;
                    const int w = s * Wp + x;
                    const uint32_t wl = __ldg(wall + w) & m, walk = ~__ldg(wall + w) & m;
                    const uint32_t worth = __ldg(tpos + w) & m;
                    const uint32_t c0 = cov[w], so = seen[w];
                    const uint32_t nf = walk & ~c0, al = walk & c0;
                    if (nf) cov[w] = c0 | nf;
                    if (wl & ~so) seen[w] = so | wl;
                    cnt_new += __popc(nf);
                    const uint32_t nfw = nf & worth;
                    rel_new |= k ? (nfw << (32 - sh)) : (nfw >> sh);
                    rel_al |= k ? (al << (32 - sh)) : (al >> sh);
                }
            }
            if (lane < S) sm.rows[i * S + lane] = make_uint2(rel_new, rel_al);
            __syncwarp();
        }
        // float64 accumulation in the reference's raster order (x rows outer, y inner)
        if (lane < N) {
            for (int r = 0; r < S; ++r) {
                const uint2 rb = sm.rows[lane * S + r];
                uint32_t both = rb.x | rb.y;
                while (both) {
                    const uint32_t bit = both & (0u - both);
                    both ^= bit;
                    if (rb.x & bit) {                                            // += grid[j][k]
                        double v = 1.0;
                        if (p.cell_values)
                            v = __ldg(p.cell_values + ((size_t)map_idx * W + (unpack_x(pv) - R + r)) * L +
                                      (unpack_y(pv) - R + (__ffs(bit) - 1)));
                        racc = __dadd_rn(racc, v);
                    } else {
                        racc = __dsub_rn(racc, p.free_penalty);                 // -= free_penalty
                    }
                }
            }
            // motion_penalty quirk (:203-208): -1 per slot on the very first step
            if (head->flags & 1u) racc = __dadd_rn(racc, -1.0);
            sm.rew[rank] = racc;
        }
        cnt_new = (int)__reduce_add_sync(FULL, cnt_new);
        __syncwarp();
        if (lane < N && pv != myv) hpos[lane] = pv;
        if (lane == 0) {
            double r = numpy_sum(sm.rew, N);                                    // :214
            const int cs = head->currstep + 1;                                  // :211
            const int covered = hcov[0] + cnt_new;
            const int fc = __ldg(p.free_counts + map_idx);
            double th = head->done_thresh;
            int dn = 0;
            bool full;                                                         // :215, :401-414
            if (th >= 1.0) full = covered >= fc;
            else full = th <= (double)covered / (double)fc;
            if (full) { r += p.terminal; th += p.done_incr; dn = 1; head->done_thresh = th; }
            const double epr = head->ep_return + r;
            head->currstep = cs;
            head->flags = head->flags & ~1u;
            head->ep_return = epr;
            if (cnt_new) hcov[0] = covered;
            slot.covered = covered;
            const bool trunc = p.truncate_steps > 0 && cs >= p.truncate_steps;
            sm.stats[4] += (double)ncoll;
            sm.stats[5] += (double)cnt_new;
            sm.stats[6] += (double)N;
            int nr = 0;
            if (dn || trunc) {
                sm.stats[0] += 1.0;
                sm.stats[1] += epr;
                sm.stats[2] += (double)cs;
                sm.stats[3] += (double)covered / (double)fc;
                nr = 1;
            }
            p.reward[e] = (float)r;
            if (p.reward64) p.reward64[e] = r;
            p.done[e] = (uint8_t)(dn | (trunc ? MCOV_DONE_TRUNCATED : 0));   // bit 1: cut at truncate_steps
            again = (nr && p.auto_reset) ? 1 : 0;
            p.needs_reset[e] = (nr && !p.auto_reset) ? 1 : 0;
        }
    } else if (lane == 0) {
        slot.covered = hcov[0];
        if (mode == MODE_STEP) {                                               // sentinel (:88-90)
            p.reward[e] = 0.0f;
            if (p.reward64) p.reward64[e] = 0.0;
            p.done[e] = 1;
            again = p.auto_reset ? 1 : 0;
            p.needs_reset[e] = p.auto_reset ? 0 : 1;
        }
    }
        if (lane < N) slot.pos[lane] = pv;
    }
    again = __shfl_sync(FULL, again, 0);
    __syncwarp();
    if (again) super_reset_env(p, slot, e, lane);             // fused auto-reset
}

// Persistent CTA, software-pipelined over its envs: while warps 1..15 stream the state of env e,
// warp 0 already resolves the dynamics of the CTA's next env into the other slot, so the serial
// part of a step (moves, footprints, float64 reward) is hidden behind the HBM-bound writes.
__global__ void __launch_bounds__(ST, MCOV_SUPER_MINB)
super_env_kernel(const SuperParams p, const int dt_cells) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    __shared__ SuperSmem sm;
    __shared__ EnvSlot slots[2];
    SuperDyn dy;
    dy.red = reinterpret_cast<int*>(smem_raw);
    const int nred = dt_scratch_ints(p.g.L, ST);
    dy.dlut = reinterpret_cast<float*>(dy.red + nred);
    const int head_bytes = ((nred + p.g.W + p.g.L + 2) * 4 + 15) & ~15;
    dy.dbuf = p.dt_scratch ? p.dt_scratch + (size_t)blockIdx.x * dt_cells
                           : reinterpret_cast<uint16_t*>(smem_raw + head_bytes);
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    if (threadIdx.x < 16) {
        const uint32_t t = threadIdx.x;
        sm.lut[t] = make_float4((t & 1u) ? 1.0f : 0.0f, (t & 2u) ? 1.0f : 0.0f, (t & 4u) ? 1.0f : 0.0f,
                                (t & 8u) ? 1.0f : 0.0f);
    }
    if (threadIdx.x < MCOV_STATS_LEN) sm.stats[threadIdx.x] = 0.0;

    const int mode = p.mode;
    auto next_env = [&](int e) -> int {                       // next env of this CTA that the call touches
        for (e += gridDim.x; e < p.E; e += gridDim.x) {
            if (mode != MODE_RESET) return e;
            if (p.mask ? p.mask[e] : (p.mask_from_flags ? p.needs_reset[e] : 1)) return e;
        }
        return -1;
    };
    // the start-of-step distance map (dist_reward, :117-118) needs the whole CTA before the moves
    auto produce = [&](int e, EnvSlot& slot) {
        if (p.dist_reward && mode == MODE_STEP) {
            const bool sentinel = p.act_fmt == MCOV_ACT_I64_BASE4 &&
                                  reinterpret_cast<const long long*>(p.actions)[e] == -1;
            if (!sentinel) super_transform(p.covered + (size_t)e * p.g.PW, p.g, dy, p.dt_pitch, sm, whole_cta());
        }
        if (warp == 0) super_dynamics(p, sm, slot, dy, e, mode, lane);
        __syncthreads();
    };

    int e = next_env((int)blockIdx.x - (int)gridDim.x);
    int s = 0;
    __syncthreads();
    if (e >= 0) produce(e, slots[0]);
    const bool overlapped = p.vec_ok && !p.dist_reward;
#ifndef MCOV_SUPER_DTW
#define MCOV_SUPER_DTW 7
#endif
    constexpr int DT_WARPS = MCOV_SUPER_DTW;                 // warps 1..DT_WARPS run the transform, the rest write
    while (e >= 0) {
        const int en = next_env(e);
        if (overlapped) {
            // warp 0: dynamics of the next env | warps 1..7: transform, then help | warps 8..15: the
            // six transform-independent planes | then warps 1..15: the distance plane
            if (threadIdx.x == 0) { sm.work = 0; sm.any_zero = 1; sm.dmax = 0; }
            __syncthreads();
            if (warp == 0) {
                if (en >= 0) super_dynamics(p, sm, slots[s ^ 1], dy, en, mode, lane);
            } else {
                if (warp <= DT_WARPS && slots[s].covered > 0) {
                    Team team;
                    team.rank = (int)threadIdx.x - 32;
                    team.size = DT_WARPS * 32;
                    team.bar = 1;
                    super_transform(p.covered + (size_t)e * p.g.PW, p.g, dy, p.dt_pitch, sm, team);
                }
                write_fixed_planes(p, sm, slots[s], e, lane);
                asm volatile("bar.sync 2, %0;" :: "r"(ST - 32) : "memory");      // transform done, seen by the writers
                write_dist_plane(p, sm, slots[s], dy, e, 32, ST - 32);
            }
            __syncthreads();
        } else {
            if (slots[s].covered > 0) {
                super_transform(p.covered + (size_t)e * p.g.PW, p.g, dy, p.dt_pitch, sm, whole_cta());
            } else {
                if (threadIdx.x == 0) { sm.any_zero = 1; sm.dmax = 0; }
                __syncthreads();
            }
            if (!p.dist_reward) {
                if (warp == 0) {
                    if (en >= 0) super_dynamics(p, sm, slots[s ^ 1], dy, en, mode, lane);
                } else {
                    write_state(p, sm, slots[s], dy, e, 32, ST - 32);
                }
                __syncthreads();
            } else {
                write_state(p, sm, slots[s], dy, e, 0, ST);
                __syncthreads();
                if (en >= 0) produce(en, slots[s ^ 1]);
            }
        }
        e = en;
        s ^= 1;
    }

    // ---- episode statistics: CTA row -> last CTA folds all rows (single launch) -------------------
    if (p.stats_out) {
        __syncthreads();
        if (threadIdx.x < MCOV_STATS_LEN)
            p.stats_partial[(size_t)blockIdx.x * MCOV_STATS_LEN + threadIdx.x] = sm.stats[threadIdx.x];
        __threadfence();
        __syncthreads();
        if (threadIdx.x == 0) sm.last = (atomicAdd(p.counter, 1u) == gridDim.x - 1) ? 1 : 0;
        __syncthreads();
        if (sm.last) {
            __threadfence();
            if (threadIdx.x < MCOV_STATS_LEN) {
                double t = 0.0;
                for (int r = 0; r < (int)gridDim.x; ++r)
                    t += __ldcg(p.stats_partial + (size_t)r * MCOV_STATS_LEN + threadIdx.x);
                p.stats_out[threadIdx.x] = t;
            }
            if (threadIdx.x == 0) *p.counter = 0u;
        }
    }
}

struct SuperShape {
    int grid, pitch, cells;
    size_t smem;                 // dynamic shared memory (0 when the work buffer is global)
    bool global_dt;
};

inline SuperShape super_shape(const mcov_env* env) {
    SuperShape s;
    s.pitch = dt_pitch_of(env->g);
    s.cells = env->g.W * s.pitch;
    const size_t bytes = (size_t)s.cells * 2;
    const size_t head = (((size_t)dt_scratch_ints(env->g.L, ST) + env->g.W + env->g.L + 2) * 4 + 15) & ~(size_t)15;
    s.global_dt = bytes + head > 200 * 1024;
    s.smem = head + (s.global_dt ? 0 : ((bytes + 15) / 16 * 16));
    int per_sm = 3;                                          // __launch_bounds__(ST, MCOV_SUPER_MINB)
    if (!s.global_dt) {
        const int by_smem = (int)((220 * 1024) / (s.smem + sizeof(SuperSmem) + 1024));
        if (per_sm > by_smem) per_sm = by_smem;
        if (per_sm < 1) per_sm = 1;
    } else {
        per_sm = 2;
    }
    int grid = env->num_sms * per_sm;
    if (grid > env->cfg.num_envs) grid = env->cfg.num_envs;
    s.grid = grid;
    return s;
}

}  // namespace

int super_grid_size(const mcov_env* env) { return super_shape(env).grid; }

int64_t super_dt_scratch_bytes(const mcov_env* env) {
    const SuperShape s = super_shape(env);
    return s.global_dt ? (int64_t)s.grid * s.cells * 2 : 0;
}

cudaError_t launch_super(mcov_env* env, SuperParams& p, cudaStream_t st) {
    const SuperShape s = super_shape(env);
    p.dt_pitch = s.pitch;
    if (s.global_dt && !p.dt_scratch) return cudaErrorInvalidValue;
    if (!s.global_dt) p.dt_scratch = nullptr;
    if (s.smem > env->super_smem_configured) {                // per handle = per device (see launch_dec_t)
        cudaError_t err = cudaSuccess;
        if (s.smem > 38 * 1024)
            err = cudaFuncSetAttribute(super_env_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)s.smem);
        if (err == cudaSuccess)
            err = cudaFuncSetAttribute(super_env_kernel, cudaFuncAttributePreferredSharedMemoryCarveout,
                                       (int)cudaSharedmemCarveoutMaxShared);
        if (err != cudaSuccess) return err;
        env->super_smem_configured = s.smem;
    }
    super_env_kernel<<<s.grid, ST, s.smem, st>>>(p, s.cells);
    env->launches++;
    return cudaGetLastError();
}

}  // namespace mcov
