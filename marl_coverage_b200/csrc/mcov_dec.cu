// mcov_dec.cu — DecGridRL reset/step/observe as ONE fused sm_100a kernel.
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
// Mapping: one CTA owns one env per loop iteration (persistent grid-stride loop); inside the CTA a
// group of LPA lanes (16 or 32) owns one robot and its lanes own window ROWS.  All per-robot maps
// are bit planes in HBM ("strip-major", see include/mcov.h); a window row is a <=32-bit field cut
// from <=2 words with a funnel shift.  Newly covered cells are counted with __popc + shuffles, no
// global atomics: every bitmap word has exactly one writer (owner election when footprints
// overlap).  The float32 observation block of the env is assembled in shared memory and leaves
// the SM as ONE async bulk copy (cp.async.bulk, TMA engine) straight into the torch buffer.
#include "mcov_internal.h"

namespace mcov {

namespace {

struct DecSmem {
    float* stage;        // [obs_env_floats]
    uint32_t* winF;      // [N][S] sensor-window free hits, window-relative bits
    uint32_t* winO;      // [N][S] sensor-window obstacle hits
    uint32_t* wallW;     // [N][S] lidar: truth wall window
    uint32_t* posF;      // [N] final positions packed (x | y<<16)
    int* covS;           // [N] per-robot known-free counts after this step
    int* red;            // [32] per-warp partial new-cell counts
    int* misc;           // [4]
    double* stats;       // [MCOV_STATS_LEN] per-CTA running sums (thread 0 only)
};

__host__ __device__ inline size_t dec_smem_bytes(int obs_env_floats, int N, int S) {
    size_t b = (size_t)round_up(obs_env_floats * 4, 16);
    b += (size_t)3 * N * S * 4;
    b += (size_t)N * 4 * 2;
    b += 32 * 4 + 4 * 4;
    b += MCOV_STATS_LEN * 8;
    return round_up((int)b, 16);
}

__device__ __forceinline__ DecSmem carve(unsigned char* base, int obs_env_floats, int N, int S) {
    DecSmem s;
    s.stage = reinterpret_cast<float*>(base);
    unsigned char* q = base + round_up(obs_env_floats * 4, 16);
    s.stats = reinterpret_cast<double*>(q); q += MCOV_STATS_LEN * 8;
    s.winF = reinterpret_cast<uint32_t*>(q); q += N * S * 4;
    s.winO = reinterpret_cast<uint32_t*>(q); q += N * S * 4;
    s.wallW = reinterpret_cast<uint32_t*>(q); q += N * S * 4;
    s.posF = reinterpret_cast<uint32_t*>(q); q += N * 4;
    s.covS = reinterpret_cast<int*>(q); q += N * 4;
    s.red = reinterpret_cast<int*>(q); q += 32 * 4;
    s.misc = reinterpret_cast<int*>(q);
    return s;
}

__device__ __forceinline__ uint32_t pack_xy(int x, int y) {
    return (uint32_t)(x & 0xffff) | ((uint32_t)(y & 0xffff) << 16);
}
__device__ __forceinline__ int unpack_x(uint32_t v) { return (int)(int16_t)(v & 0xffffu); }
__device__ __forceinline__ int unpack_y(uint32_t v) { return (int)(int16_t)(v >> 16); }

// Draws map + start poses on the device: uniform map, uniform distinct walkable cells by rejection
// (same distribution as dec_grid_rl.py:462-502; not numpy's Mersenne-Twister stream).
__device__ void draw_reset(const DecParams& p, int e, uint32_t episode, uint32_t* posS, int* map_out) {
    const uint2 key = make_uint2((uint32_t)p.seed, (uint32_t)(p.seed >> 32));
    const uint32_t gid = (uint32_t)(p.env_id_base + e);
    int mi;
    if (p.map_idx_in) {
        mi = p.map_idx_in[e];
    } else {
        const uint4 r = philox4x32_10(make_uint4(gid, episode, 0u, 0x4D434F56u), key);
        mi = bounded(r.x, p.n_maps);
    }
    *map_out = mi;
    if (p.start_xy_in) {
        const int16_t* s = p.start_xy_in + (size_t)e * p.N * 2;
        for (int j = 0; j < p.N; ++j) posS[j] = pack_xy(s[2 * j], s[2 * j + 1]);
        return;
    }
    const uint32_t* wall = p.wall + (size_t)mi * p.g.PW;
    int placed = 0;
    for (uint32_t k = 1; placed < p.N && k < 8192u; ++k) {
        const uint4 r = philox4x32_10(make_uint4(gid, episode, k, 0x4D434F56u), key);
        const uint32_t cand[4] = {r.x, r.y, r.z, r.w};
        for (int h = 0; h < 2 && placed < p.N; ++h) {
            const int x = bounded(cand[2 * h], p.g.W), y = bounded(cand[2 * h + 1], p.g.L);
            if (get_bit(wall, p.g.Wp, x, y)) continue;
            bool taken = false;
            for (int j = 0; j < placed; ++j) taken |= (posS[j] == pack_xy(x, y));
            if (!taken) posS[placed++] = pack_xy(x, y);
        }
    }
    // pathological map (almost no walkable cell): deterministic scan so the kernel always ends
    for (int x = 0; x < p.g.W && placed < p.N; ++x)
        for (int y = 0; y < p.g.L && placed < p.N; ++y) {
            if (get_bit(wall, p.g.Wp, x, y)) continue;
            bool taken = false;
            for (int j = 0; j < placed; ++j) taken |= (posS[j] == pack_xy(x, y));
            if (!taken) posS[placed++] = pack_xy(x, y);
        }
    for (; placed < p.N; ++placed) posS[placed] = pack_xy(0, 0);
}

template <int LPA>
__global__ void __launch_bounds__(LPA == 16 ? 256 : 1024, LPA == 16 ? 4 : 1)
dec_env_kernel(const DecParams p) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    const DecSmem sm = carve(smem_raw, p.obs_env_floats, p.N, p.S);

    const int tid = threadIdx.x;
    const int lane = tid & 31;
    const int warp = tid >> 5;
    const int nwarps = (blockDim.x + 31) >> 5;
    const int lg = tid % LPA;             // lane within the robot group
    const int i = tid / LPA;              // robot owned by this group
    const int gshift = lane & ~(LPA - 1); // first warp lane of this group
    const bool gvalid = i < p.N;          // blockDim is rounded up to whole warps
    const uint32_t gbits = LPA == 32 ? 0xffffffffu : 0xffffu;
    const int N = p.N, S = p.S, R = p.R, W = p.g.W, L = p.g.L, Wp = p.g.Wp, strips = p.g.strips;
    const int ow = p.ow, ego = p.ego, C = p.C;
    const int mode = p.mode;

    double* st_acc = sm.stats;
    if (tid == 0)
        for (int k = 0; k < MCOV_STATS_LEN; ++k) st_acc[k] = 0.0;

    for (int e = blockIdx.x; e < p.E; e += gridDim.x) {
        unsigned char* hrec = p.hdr + (size_t)e * p.hdr_stride;
        EnvHead* head = reinterpret_cast<EnvHead*>(hrec);
        uint32_t* hpos = reinterpret_cast<uint32_t*>(hrec + p.off_pos);
        int32_t* hcov = reinterpret_cast<int32_t*>(hrec + p.off_cov);
        uint32_t* hadj = reinterpret_cast<uint32_t*>(hrec + p.off_adj);
        uint2* fo_i = p.fo + ((size_t)e * N + (gvalid ? i : 0)) * p.g.PW;
        uint32_t* vis = p.visited + (size_t)e * p.g.PW;

        if (mode == MODE_RESET) {
            int act = 1;
            if (p.mask) act = p.mask[e];
            else if (p.mask_from_flags) act = p.needs_reset[e];
            if (!act) continue;                                   // CTA-uniform
        }

        // ------------------------------------------------------------------ phase A
        const bool fresh = (mode == MODE_RESET);
        int map_idx;
        uint32_t shown = 0;
        if (fresh) {
            if (tid == 0) bulk_wait_read_all();                   // stage/posF reuse
            if (tid == 0) {
                int mi;
                draw_reset(p, e, head->episode, sm.posF, &mi);
                sm.misc[0] = mi;
            }
            // wipe this env's planes (dec_grid_rl.py:505-514) with 128-bit stores
            {
                uint4* z = reinterpret_cast<uint4*>(p.fo + (size_t)e * N * p.g.PW);
                const int n16 = N * p.g.PW / 2;
                for (int k = tid; k < n16; k += blockDim.x) z[k] = make_uint4(0, 0, 0, 0);
                uint4* zv = reinterpret_cast<uint4*>(vis);
                for (int k = tid; k < p.g.PW / 4; k += blockDim.x) zv[k] = make_uint4(0, 0, 0, 0);
            }
            __syncthreads();
            map_idx = sm.misc[0];
        } else {
            map_idx = head->map_idx;
            shown = head->flags;
        }
        const uint32_t* wall = p.wall + (size_t)map_idx * p.g.PW;
        const uint32_t* tpos = p.pos + (size_t)map_idx * p.g.PW;

        int px = 0, py = 0;
        if (lg < N) {
            const uint32_t v = fresh ? sm.posF[lg] : hpos[lg];
            px = unpack_x(v);
            py = unpack_y(v);
        }

        int ncoll = 0;
        bool sentinel = false;
        if (mode == MODE_STEP) {
            // ---- action decode (dec_grid_rl.py:110-117) ------------------------------------
            int u = 0;
            if (p.act_fmt == MCOV_ACT_I64_BASE4) {
                const long long a = reinterpret_cast<const long long*>(p.actions)[e];
                sentinel = (a == -1);
                u = (int)((a >> (2 * (lg & 31))) & 3);
            } else if (lg < N) {
                u = reinterpret_cast<const uint8_t*>(p.actions)[(size_t)e * N + lg] & 3;
            }
            if (!sentinel) {
                // ---- sequential, collision-checked moves (dec_grid_rl.py:128-145,171-204) ----
                // Dec table: 0:x+1  1:y+1  2:x-1  3:y-1
                const int tx = px + (u == 0) - (u == 2);
                const int ty = py + (u == 1) - (u == 3);
                int blocked0 = 1;
                if (lg < N && tx >= 0 && tx < W && ty >= 0 && ty < L)
                    blocked0 = get_bit(wall, Wp, tx, ty);
                for (int k = 0; k < N; ++k) {
                    const int txk = __shfl_sync(0xffffffffu, tx, k, LPA);
                    const int tyk = __shfl_sync(0xffffffffu, ty, k, LPA);
                    const int blk = __shfl_sync(0xffffffffu, blocked0, k, LPA);
                    const bool hit = (lg < N) && px == txk && py == tyk;
                    const uint32_t b = (__ballot_sync(0xffffffffu, hit) >> gshift) & gbits;
                    if (!blk && b == 0) {
                        if (lg == k) { px = txk; py = tyk; }
                        shown |= 1u << k;
                    } else {
                        ++ncoll;
                    }
                }
            }
        }
        // publish poses for cross-robot lookups; group 0 also persists them
        if (i == 0 && lg < N) {
            const uint32_t v = pack_xy(px, py);
            sm.posF[lg] = v;
            if (mode != MODE_OBSERVE) hpos[lg] = v;
        }
        const int xi = __shfl_sync(0xffffffffu, px, i, LPA);
        const int yi = __shfl_sync(0xffffffffu, py, i, LPA);

        // ---- communication graph (dec_grid_rl.py:374-391): Chebyshev ball, diagonal included --
        if (p.want_adj && mode != MODE_OBSERVE) {
            const bool link = (lg < N) && max(abs(px - xi), abs(py - yi)) <= p.comm_radius;
            const uint32_t b = (__ballot_sync(0xffffffffu, link) >> gshift) & gbits;
            if (lg == 0 && gvalid) hadj[i] = b;
        }

        const bool do_sense = (mode == MODE_RESET) || (mode == MODE_STEP && !sentinel);
        bool near_other = false;
        if (do_sense && gvalid) {
            // can another robot's sensor rows share a bitmap word with mine?
            near_other = (lg < N) && (lg != i) && abs(px - xi) <= 2 * R && abs(py - yi) < S + 32;
            const int y0 = yi - R;
            if (p.sensor == MCOV_SENSOR_SQUARE) {
                // squaresensor.py:15-37 — the footprint is the truth map itself (sees through walls)
                for (int r = lg; r < S; r += LPA) {
                    const int x = xi - R + r;
                    uint32_t f = 0, o = 0;
                    if (x >= 0 && x < W) {
                        o = read_window(wall, Wp, strips, x, y0, S);
                        if (p.single_tool) f = (r == R) ? (1u << R) : 0u;      // dec_grid_rl.py:233-234
                        else f = read_window(tpos, Wp, strips, x, y0, S);
                    }
                    sm.winF[i * S + r] = f;
                    sm.winO[i * S + r] = o;
                }
            } else {
                // lidar.py:16-65 — stage the truth wall window, then march beams in float64
                for (int r = lg; r < S; r += LPA) {
                    const int x = xi - R + r;
                    sm.wallW[i * S + r] = (x >= 0 && x < W) ? read_window(wall, Wp, strips, x, y0, S)
                                                            : 0xffffffffu;
                    sm.winF[i * S + r] = 0;
                    sm.winO[i * S + r] = 0;
                }
                __syncwarp();
                const double rng = (double)R;
                for (int l = lg; l < p.n_lasers; l += LPA) {
                    const double xinc = p.beams[3 * l], yinc = p.beams[3 * l + 1], dinc = p.beams[3 * l + 2];
                    double cx = (double)xi, cy = (double)yi, dist = 0.0;
                    int rx, ry;
                    bool walk;
                    while (true) {
                        const bool inside = cx >= 0.0 && cy >= 0.0 && cx < (double)W && cy < (double)L;
                        rx = __double2int_rz(cx) - (xi - R);
                        ry = __double2int_rz(cy) - y0;
                        const bool inwin = rx >= 0 && rx < S && ry >= 0 && ry < S;
                        walk = inside && inwin && !((sm.wallW[i * S + rx] >> ry) & 1u);
                        if (!(walk && dist < rng)) break;
                        atomicOr(&sm.winF[i * S + rx], 1u << ry);
                        cx += xinc;
                        cy += yinc;
                        dist += dinc;
                    }
                    if (rx >= 0 && rx < S && ry >= 0 && ry < S) {
                        if (walk) atomicOr(&sm.winF[i * S + rx], 1u << ry);
                        else atomicOr(&sm.winO[i * S + rx], 1u << ry);
                    }
                }
                if (p.single_tool) {
                    __syncwarp();
                    for (int r = lg; r < S; r += LPA) sm.winF[i * S + r] = (r == R) ? (1u << R) : 0u;
                }
            }
        }
        if (tid == 0 && !fresh) bulk_wait_read_all();             // previous env's obs left smem
        const int overlap = __syncthreads_or(near_other ? 1 : 0); // sync #1: windows + poses published

        // ------------------------------------------------------------------ phase B
        int cnt = 0, covnew = 0;
        if (do_sense && gvalid) {
            const int y0 = yi - R;
            const int s0 = y0 >> 5;
            const bool straddle = (y0 & 31) + S > 32;
            for (int r = lg; r < S; r += LPA) {
                const int x = xi - R + r;
                if (x < 0 || x >= W) continue;
                uint32_t fw[2], owd[2];
                split_words(sm.winF[i * S + r], y0, fw[0], fw[1]);
                split_words(sm.winO[i * S + r], y0, owd[0], owd[1]);
#pragma unroll
                for (int k = 0; k < 2; ++k) {
                    const int s = s0 + k;
                    if (s < 0 || s >= strips || (k == 1 && !straddle)) continue;
                    const int widx = s * Wp + x;
                    // per-robot memory: free_i |= hits, obst_i |= hits  (dec_grid_rl.py:231-236)
                    const uint2 old = fresh ? make_uint2(0, 0) : fo_i[widx];
                    const uint32_t nf = old.x | fw[k], no = old.y | owd[k];
                    covnew += __popc(fw[k] & ~old.x);
                    if (nf != old.x || no != old.y) fo_i[widx] = make_uint2(nf, no);
                    // union "visited" plane: exactly one writer per word (dec_grid_rl.py:243-256)
                    uint32_t comb = fw[k];
                    bool owner = true;
                    if (overlap) {
                        for (int j = 0; j < N; ++j) {
                            if (j == i) continue;
                            const uint32_t pj = sm.posF[j];
                            const int rj = x - (unpack_x(pj) - R);
                            if (rj < 0 || rj >= S) continue;
                            const int y0j = unpack_y(pj) - R;
                            const int s0j = y0j >> 5;
                            uint32_t lo, hi;
                            split_words(sm.winF[j * S + rj], y0j, lo, hi);
                            if (s == s0j) { comb |= lo; owner &= (j > i); }
                            else if (s == s0j + 1 && (y0j & 31) + S > 32) { comb |= hi; owner &= (j > i); }
                        }
                    }
                    if (owner) {
                        const uint32_t vold = fresh ? 0u : vis[widx];
                        const uint32_t nb = comb & ~vold;
                        cnt += __popc(nb);
                        if (nb) vis[widx] = vold | comb;
                    }
                }
            }
        }
        // per-robot known-free count (percent_covered sums over robots, quirk A.3.3)
#pragma unroll
        for (int d = LPA / 2; d > 0; d >>= 1) covnew += __shfl_xor_sync(0xffffffffu, covnew, d, LPA);
        if (lg == 0 && gvalid) {
            const int c = (fresh ? 0 : hcov[i]) + covnew;
            sm.covS[i] = c;
            if (covnew != 0 || fresh) hcov[i] = c;
        }
#pragma unroll
        for (int d = 16; d > 0; d >>= 1) cnt += __shfl_xor_sync(0xffffffffu, cnt, d);
        if (lane == 0) sm.red[warp] = cnt;

        // ---- egocentric observation rows (dec_grid_rl.py:334-347) -> float32 staging -----------
        if (C > 3 && gvalid) {
            float* z = sm.stage + (size_t)(i * C + 3) * ow * ow;
            for (int k = lg; k < (C - 3) * ow * ow; k += LPA) z[k] = 0.0f;
        }
        if (gvalid) {
            const int y0e = yi - ego;
            const uint32_t wm = low_mask(ow);
            for (int a = lg; a < ow; a += LPA) {
                const int x = xi - ego + a;
                uint32_t bR = 0, bF = 0, bO = 0;
                if (x >= 0 && x < W) {
                    if (!fresh) {
                        const uint2 w2 = read_window2(fo_i, Wp, strips, x, y0e, ow);
                        bF = w2.x;
                        bO = w2.y;
                    }
                    if (do_sense) {
                        const int r = a - (ego - R);
                        if (r >= 0 && r < S) {
                            bF |= shift_bits(sm.winF[i * S + r], ego - R) & wm;
                            bO |= shift_bits(sm.winO[i * S + r], ego - R) & wm;
                        }
                    }
                    // layer 0 shows every robot that has moved at least once (quirk A.3.4)
                    for (int j = 0; j < N; ++j) {
                        if (!((shown >> j) & 1u)) continue;
                        const uint32_t pj = sm.posF[j];
                        const int d = unpack_y(pj) - y0e;
                        if (unpack_x(pj) == x && d >= 0 && d < ow) bR |= 1u << d;
                    }
                }
                float* row = sm.stage + ((size_t)(i * C) * ow + a) * ow;
                for (int b = 0; b < ow; ++b) {
                    row[b] = ((bR >> b) & 1u) ? 1.0f : 0.0f;
                    row[ow * ow + b] = ((bF >> b) & 1u) ? 1.0f : 0.0f;
                    row[2 * ow * ow + b] = ((bO >> b) & 1u) ? 1.0f : 0.0f;
                }
            }
        }
        fence_proxy_async_smem();
        __syncthreads();                                          // sync #2: stage + partials ready

        // ------------------------------------------------------------------ phase C
        float* obs_e = p.obs + (size_t)e * p.obs_env_floats;
        if (!p.bulk_ok) {
            for (int k = tid; k < p.obs_env_floats; k += blockDim.x) obs_e[k] = sm.stage[k];
        }
        if (tid == 0) {
            if (p.bulk_ok) bulk_store_s2g(obs_e, sm.stage, (uint32_t)p.obs_env_floats * 4u);
            if (mode == MODE_STEP) {
                double r = 0.0;
                int dn = 0;
                if (sentinel) {
                    dn = 1;                                        // dec_grid_rl.py:105-107
                } else {
                    for (int k = 0; k < ncoll; ++k) r -= p.cp;     // dec_grid_rl.py:185-204
                    int newc = 0;
                    for (int w = 0; w < nwarps; ++w) newc += sm.red[w];
                    double obs_r = (double)newc;
                    if (p.dist_term) obs_r = (double)p.dist_term[e] + (double)newc;   // row D11
                    r += obs_r;
                    const int cs = head->currstep + 1;             // :154
                    int covsum = 0;
                    for (int j = 0; j < N; ++j) covsum += sm.covS[j];
                    const double pct = (double)covsum / (double)p.free_counts[map_idx];
                    double th = head->done_thresh;
                    const bool full = fmin(th, 1.0) <= pct;        // :156-157, :540-546
                    if (full) { r += p.terminal; th += p.done_incr; dn = 1; }
                    else if (cs == p.maxsteps) dn = 1;
                    const double epr = head->ep_return + r;
                    head->currstep = cs;
                    head->flags = shown;
                    head->done_thresh = th;
                    head->ep_return = epr;
                    const bool trunc = p.truncate_steps > 0 && cs >= p.truncate_steps;
                    st_acc[4] += (double)ncoll;
                    st_acc[5] += (double)newc;
                    st_acc[6] += (double)N;
                    if (dn || trunc) {
                        st_acc[0] += 1.0;
                        st_acc[1] += epr;
                        st_acc[2] += (double)cs;
                        st_acc[3] += pct;
                    }
                    p.needs_reset[e] = (dn || trunc) ? 1 : 0;
                }
                if (sentinel) p.needs_reset[e] = 1;
                p.reward[e] = (float)r;
                if (p.reward64) p.reward64[e] = r;
                p.done[e] = (uint8_t)dn;
            } else if (mode == MODE_RESET) {
                head->currstep = 0;                                // :479
                head->map_idx = map_idx;
                head->flags = 0;                                   // _robot_pad starts empty (:487)
                head->episode = head->episode + 1;
                head->ep_return = 0.0;
                p.needs_reset[e] = 0;
            }
        }
    }
    if (tid == 0) {
        bulk_wait_all();
        if (p.stats_partial) {
#pragma unroll
            for (int k = 0; k < MCOV_STATS_LEN; ++k)
                p.stats_partial[(size_t)blockIdx.x * MCOV_STATS_LEN + k] = st_acc[k];
        }
    }
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

inline int pick_lpa(const mcov_config& c) {
    const int S = 2 * c.sensor_range + 1, ow = 2 * c.egoradius + 1;
    const bool small = c.num_robots <= 16 && S <= 16 && ow <= 16 && c.sensor_type == MCOV_SENSOR_SQUARE;
    return small ? 16 : 32;
}

}  // namespace

int dec_grid_size(const mcov_env* env) {
    const int lpa = pick_lpa(env->cfg);
    const int threads = env->cfg.num_robots * lpa;
    const int S = 2 * env->cfg.sensor_range + 1, ow = 2 * env->cfg.egoradius + 1;
    const size_t smem = dec_smem_bytes(env->cfg.num_robots * env->obs_C * ow * ow, env->cfg.num_robots, S);
    int per_sm = 2048 / round_up(threads, 32);
    if (per_sm > 32) per_sm = 32;
    const int by_smem = (int)((220 * 1024) / (smem + 1024));
    if (per_sm > by_smem) per_sm = by_smem;
    if (per_sm < 1) per_sm = 1;
    int grid = env->num_sms * per_sm;
    if (grid > env->cfg.num_envs) grid = env->cfg.num_envs;
    return grid;
}

cudaError_t launch_dec(mcov_env* env, DecParams& p, cudaStream_t st) {
    const int lpa = pick_lpa(env->cfg);
    const int threads = round_up(p.N * lpa, 32);
    const size_t smem = dec_smem_bytes(p.obs_env_floats, p.N, p.S);
    const int grid = dec_grid_size(env);
    cudaError_t err;
    if (lpa == 16) {
        if (smem > 48 * 1024) {
            err = cudaFuncSetAttribute(dec_env_kernel<16>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
            if (err != cudaSuccess) return err;
        }
        dec_env_kernel<16><<<grid, threads, smem, st>>>(p);
    } else {
        if (smem > 48 * 1024) {
            err = cudaFuncSetAttribute(dec_env_kernel<32>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
            if (err != cudaSuccess) return err;
        }
        dec_env_kernel<32><<<grid, threads, smem, st>>>(p);
    }
    env->launches++;
    return cudaGetLastError();
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
