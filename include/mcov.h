/*
 * mcov.h — C ABI of libmcov_b200.so: batched MARL-Coverage environment step for NVIDIA B200 (sm_100a).
 *
 * This is the drop-in boundary for ONE hot path of ExistentialRobotics/MARL-Coverage: the
 * reset()/step(actions) path of its two grid environments, their sensors and discrete moves.
 * Every entry point names the reference interface it replaces (file:line relative to the
 * reference root).  The reference is pure Python and has no FFI of its own; the binding a
 * maintainer adds is a ctypes stub — see INTEGRATION.md — and marl_coverage_b200/_lib.py is
 * exactly that stub.
 *
 * Conventions
 *   - plain C types only; every pointer suffixed _dev is DEVICE memory owned by the caller
 *     (PyTorch allocates it), every pointer suffixed _host is host memory;
 *   - `stream` is a cudaStream_t passed as void*; all work is enqueued asynchronously on it,
 *     nothing synchronises, everything is CUDA-graph capturable (tests/test_graph_gpu.py; scalar arguments
 *     such as the reset seed are frozen into a captured graph like any kernel parameter);
 *   - every function returns MCOV_OK (0) or a negative mcov_status; the message is available from
 *     mcov_last_error() (thread local).  Nothing throws, nothing calls exit();
 *   - the library owns only the opaque handle.  Simulator state lives in a caller-owned arena
 *     (mcov_state_bytes / mcov_bind_state) whose layout is reported by mcov_state_layout so the
 *     host side can view/patch any field (get_state / set_state) without further entry points;
 *   - one handle <-> one device (the one current at mcov_create; a call with another device current returns
 *     MCOV_ERR_STATE); calls on one handle must be serialised by the caller;
 *   - values read from caller-supplied DEVICE buffers that become addresses (map indices, start poses, replay
 *     sample indices) are clamped into range by the kernels; they are never trusted.
 *
 * Bit-plane layout ("strip-major"): a W x L boolean plane is stored as 32-bit words,
 *   word index = strip * Wp + x,   bit b of that word = cell (x, y = 32*strip + b),
 *   strips = ceil(L / 32),  Wp = W rounded up to a multiple of 4 (16-byte strip pitch),
 * so the (2r+1)^2 window around an agent is <= 2 short contiguous word runs.  Axis 0 of every
 * reference array is x, axis 1 is y.
 */
#ifndef MCOV_H_
#define MCOV_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define MCOV_ABI_VERSION 8
#define MCOV_MAX_ROBOTS 32       /* base-4 joint actions live in an int64: 31 digits; lanes: 32 */
#define MCOV_MAX_RADIUS 15       /* sensor range, egoradius, mini_map_rad: windows fit 32 bits  */
#define MCOV_MAX_LASERS 1024
#define MCOV_STATS_LEN 8
/* done_out bits of mcov_step */
#define MCOV_DONE_TERMINAL 1     /* the reference's done(): coverage threshold reached / Dec maxsteps / sentinel */
#define MCOV_DONE_TRUNCATED 2    /* episode cut at truncate_steps (generate_episode's done = True, utils.py:25-31) */

typedef struct mcov_env mcov_env; /* opaque */

typedef enum mcov_status {
    MCOV_OK = 0,
    MCOV_ERR_BAD_CONFIG = -1,   /* value out of range / unsupported combination               */
    MCOV_ERR_NULL = -2,         /* required pointer is NULL                                     */
    MCOV_ERR_ALIGN = -3,        /* buffer not aligned as documented                             */
    MCOV_ERR_CUDA = -4,         /* a CUDA runtime call failed (message has the CUDA error)      */
    MCOV_ERR_STATE = -5,        /* call order violated (maps / arena not bound yet)             */
    MCOV_ERR_ABI = -6           /* struct_size does not match this library                      */
} mcov_status;

enum { MCOV_ENV_DEC = 0, MCOV_ENV_SUPER = 1 };
enum { MCOV_SENSOR_SQUARE = 0, MCOV_SENSOR_LIDAR = 1 };
enum { MCOV_ACT_U8_PER_ROBOT = 0,   /* uint8  [E][N], values 0..3                              */
       MCOV_ACT_I64_BASE4 = 1 };    /* int64  [E], digit i (base 4, LSD first) = robot i;
                                       -1 = the reference's "no-op, report done" sentinel       */

/* Mirrors the reference's env_config dictionaries key by key:
 *   DecGridRL.__init__   Environments/dec_grid_rl.py:49-75
 *   SuperGridRL.__init__ Environments/super_grid_rl.py:44-54                                    */
typedef struct mcov_config {
    int32_t struct_size;        /* = sizeof(mcov_config)                                        */
    int32_t env_type;           /* MCOV_ENV_DEC | MCOV_ENV_SUPER                                */
    int32_t num_envs;           /* E: envs held by this handle (this GPU's shard)               */
    int32_t num_robots;         /* numrobot                                                     */
    int32_t env_id_base;        /* global id of env 0 (multi-GPU shards; keys the device RNG)   */
    int32_t maxsteps;           /* Dec: done() also fires at _currstep == maxsteps (:544); <=0 off */
    int32_t truncate_steps;     /* batched extension: flag needs_reset at this step count
                                   (utils.py:25-28 train/test_maxsteps); <=0 off                */
    int32_t auto_reset;         /* batched extension: step() re-seeds finished envs in place    */
    double collision_penalty;
    double terminal_reward;
    double done_thresh;         /* initial value; per-env state afterwards (curriculum)         */
    double done_incr;
    double free_penalty;        /* Super                                                        */
    int32_t dist_reward;        /* both                                                         */
    /* Dec only */
    int32_t sensor_type;        /* 'square_sensor' | 'lidar' (dec_grid_rl.py:71-75)             */
    int32_t sensor_range;       /* sensor_config['range']                                       */
    int32_t num_lasers;         /* sensor_config['num_lasers'] (odd, lidar.py:11)               */
    int32_t egoradius;
    int32_t mini_map_rad;
    int32_t comm_radius;
    int32_t allow_comm;
    int32_t map_sharing;
    int32_t single_square_tool;
    int32_t dijkstra_input;
    /* Super only */
    int32_t senseradius;
    int32_t use_scanning;
    int32_t reserved[5];
} mcov_config;

/* Byte offsets into the state arena (all arrays are env-major, see DESIGN.md "HBM layout"). */
typedef struct mcov_layout {
    int32_t struct_size;
    int32_t W, L;               /* env-coordinate map size (Dec: bordered, +2)                  */
    int32_t strips, Wp;         /* plane geometry: plane_words = strips * Wp                    */
    int32_t plane_words;
    int32_t hdr_stride;         /* bytes per env header record                                  */
    int32_t hdr_off_pos;        /* within the record: int16 (x, y) [N]                          */
    int32_t hdr_off_cov;        /* int32 per-robot known-free counts [N] (Dec) / covered count  */
    int32_t hdr_off_adj;        /* uint32 adjacency row bitmasks [N] (Dec)                      */
    /* record head: f64 done_thresh, f64 ep_return, i32 currstep, i32 map_idx, u32 shown_mask
       (Dec) / first_step_pending (Super), u32 episode                                          */
    int64_t off_hdr;            /* [E] records                                                  */
    int64_t off_visited;        /* Dec: uint32 [E][plane_words] union of known-free             */
    int64_t off_fo;             /* Dec: uint32 [E][N][plane_words][2]  (.x free_i, .y obst_i)   */
    int64_t off_covered;        /* Super: uint32 [E][plane_words]  (reference _free == 0)       */
    int64_t off_seen_obst;      /* Super: uint32 [E][plane_words]  (_observed_obstacles)        */
    int64_t off_needs_reset;    /* uint8 [E] set by step when an env finished / truncated       */
    int64_t off_scratch;        /* library scratch (stats partials, distance maps)              */
    int64_t total_bytes;
} mcov_layout;

int mcov_abi_version(void);
const char* mcov_last_error(void);

/* Plane geometry helper for the host-side packer/unpacker. */
int mcov_plane_geometry(int W, int L, int* strips_out, int* wp_out, int* words_out);

/* ctor half of DecGridRL.__init__ / SuperGridRL.__init__ (dec_grid_rl.py:30-89,
 * super_grid_rl.py:27-72): validates and stores the config.  No pygame window. */
int mcov_create(const mcov_config* cfg, mcov_env** out);
int mcov_destroy(mcov_env* env);

/* LidarSensor.__init__ (Environments/Sensors/lidar.py:6-14,39-48): per-laser float64
 * (xinc, yinc, distinc) triples computed on the host with numpy, so the marching arithmetic is
 * bit-identical to the reference's. */
int mcov_set_lidar_beams(mcov_env* env, const double* beams_host, int num_lasers);

/* Map supply (Utils/gridmaker.py:7-139 -> env ctor `train_set`): packs int8 cell grids
 * (<0 obstacle, 0 walkable-uncounted, >0 free) [n][W][L] into two strip-major planes per map
 * (wall = cell<0, pos = cell>0) plus the per-map count of cell>0; add_border=1 surrounds each map
 * with the 1-cell obstacle ring DecGridRL.reset adds (dec_grid_rl.py:471-472) and the output
 * geometry is (W+2, L+2). */
int mcov_pack_maps(const int8_t* cells_dev, int n_maps, int W, int L, int add_border,
                   uint32_t* wall_out_dev, uint32_t* pos_out_dev, int32_t* free_counts_out_dev,
                   void* stream);

/* gridgen on the device (Utils/gridmaker.py:107-139): n_maps Bernoulli maps of W x L cells, every cell an
 * obstacle with probability prob_obst and free (+1) otherwise, written straight into the packed planes
 * mcov_pack_maps would produce for them (same add_border ring, same free counts) — no host grids, no copy.
 * Same distribution as the reference's np.random.choice(a=[1.0, -1.0], p=[1 - p, p]), not numpy's stream:
 * Philox4x32-10 keyed by `seed`, counter = (map, x, y / 4), so a map does not depend on n_maps or on the
 * plane layout. */
int mcov_gen_maps(int n_maps, int W, int L, int add_border, double prob_obst, uint64_t seed,
                  uint32_t* wall_out_dev, uint32_t* pos_out_dev, int32_t* free_counts_out_dev, void* stream);

/* Binds the packed map table (kept by reference; caller keeps it alive).  W, L are the
 * env-coordinate sizes (bordered for Dec). */
int mcov_set_maps(mcov_env* env, const uint32_t* wall_dev, const uint32_t* pos_dev,
                  const int32_t* free_counts_dev, int n_maps, int W, int L);

/* Per-cell reward of covering a free cell for SuperGridRL (super_grid_rl.py:189 `reward += self._grid[j][k]`,
 * sim_env.py:159): float64 [n_maps][W][L] in map coordinates, kept by reference; NULL (the default after every
 * mcov_set_maps) = every free cell is worth 1.0, which is what the +-1 PNG maps give.  The sign planes of
 * mcov_pack_maps still decide obstacle / walkable / counted.  DecGridRL handles are refused: its reward is the
 * count of newly visited cells (dec_grid_rl.py:243-256) and never reads a cell's value. */
int mcov_set_cell_values(mcov_env* env, const double* values_dev);

/* State arena: size/layout after mcov_set_maps; the caller allocates (256-byte aligned) and binds. */
int mcov_state_layout(const mcov_env* env, mcov_layout* out);
int mcov_bind_state(mcov_env* env, void* arena_dev, size_t bytes, void* stream);

/* Observation geometry: Dec -> [E, N, C, 2e+1, 2e+1], Super -> [E, N+3 | 4, W, L]; float32. */
int mcov_obs_shape(const mcov_env* env, int32_t dims_out[5], int32_t* ndims_out);

/* DecGridRL.reset / SuperGridRL.reset (dec_grid_rl.py:449-531, super_grid_rl.py:343-399) for the
 * envs selected by mask_dev (uint8 [E], NULL = all).  map_idx_dev (int32 [E]) and start_xy_dev
 * (int16 [E][N][2], env coordinates) inject the draw; NULL -> device Philox keyed by
 * (seed, global env id, episode).  Writes the first observation of each reset env. */
int mcov_reset(mcov_env* env, const uint8_t* mask_dev, const int32_t* map_idx_dev,
               const int16_t* start_xy_dev, uint64_t seed, float* obs_out_dev, void* stream);

/* DecGridRL.step / SuperGridRL.step (dec_grid_rl.py:91-169, super_grid_rl.py:74-225) for all E
 * envs: moves with collision checks, sensing, reward, termination, observation.
 *   obs_out_dev     float32, shape mcov_obs_shape, 16-byte aligned
 *   reward_out_dev  float32 [E];  reward64_out_dev float64 [E] or NULL (reference precision)
 *   done_out_dev    uint8 [E]: bit 0 (MCOV_DONE_TERMINAL) = the reference's done flag; bit 1
 *                   (MCOV_DONE_TRUNCATED) = the episode was cut at truncate_steps this step.  With auto_reset
 *                   the env restarts in place in the same call and obs_out holds the NEW episode's first
 *                   observation: a non-zero done byte is the only sign of the episode boundary, so a
 *                   learner must treat any non-zero value as "do not bootstrap across this step"
 *   stats_out_dev   float64 [MCOV_STATS_LEN] or NULL: this step's sums over the E envs of
 *                   {episodes finished, their returns, their lengths, their final coverage,
 *                    collisions, newly covered cells, agent-steps, 0}                            */
int mcov_step(mcov_env* env, const void* actions_dev, int action_format, float* obs_out_dev,
              float* reward_out_dev, double* reward64_out_dev, uint8_t* done_out_dev,
              double* stats_out_dev, void* stream);

/* mcov_step with its launch scratch in a static device array instead of the state arena.  NO COUNTERPART IN THE
 * REFERENCE: added for callers that cannot let a library near the allocator
 * on the step path (code under CUDA graph capture, tight serving loops, harnesses that forbid cudaMalloc).
 * Same arguments, outputs and bits as mcov_step.  It calls no cudaMalloc / cudaFree or allocating routine: the
 * per-CTA episode-statistics rows and the CTA arrival counter of the launch live in a file-scope
 * `static __device__` array of MCOV_STATIC_SCRATCH_MAX_ROWS * MCOV_STATS_LEN * 8 + 256 bytes (128 KiB + 256 B) that every
 * GPU loading the library holds until the process exits; its device address is resolved with cudaGetSymbolAddress
 * once per device, on the first call with that device current (make that first call outside a graph capture).
 * Limits, each refused with a message: DecGridRL handles without map_sharing / dist_reward / dijkstra_input /
 * mini_map_rad (those steps take further launches with arena scratch) -> MCOV_ERR_BAD_CONFIG; a batch whose step
 * launch can use more than MCOV_STATIC_SCRATCH_MAX_ROWS CTAs -> MCOV_ERR_BAD_CONFIG; a handle of another device than
 * the current one -> MCOV_ERR_STATE.
 * The array is shared by EVERY call on a device, whatever the handle or stream, and keeps the last call's contents:
 * the entry point clears what must start at zero (the counter) on `stream` before the launch, the rows are
 * overwritten before they are read.  Calls that overlap on the GPU (two streams, two handles, two threads) RACE
 * on it: serialise them, or use mcov_step, whose scratch is per handle. */
#define MCOV_STATIC_SCRATCH_MAX_ROWS 2048
int mcov_step_static_scratch(mcov_env* env, const void* actions_dev, int action_format, float* obs_out_dev,
                             float* reward_out_dev, double* reward64_out_dev, uint8_t* done_out_dev,
                             double* stats_out_dev, void* stream);

/* get_egocentric_observations / get_state without stepping (dec_grid_rl.py:312-372,
 * super_grid_rl.py:305-317). */
int mcov_observe(mcov_env* env, float* obs_out_dev, void* stream);

/* percent_covered() (dec_grid_rl.py:548-552, super_grid_rl.py:416-421): float64 [E]. */
int mcov_percent_covered(mcov_env* env, double* out_dev, void* stream);

/* Replay-buffer rows for batched rollouts (the consumer right after the env step).
 * ReplayBuffer.addtransition, Policies/replaybuffer.py:33-48, for the E transitions of one batched
 * step: row (ptr + e) % capacity of the caller-owned float32 ring arrays
 *   ring_state / ring_next [capacity][row_floats], ring_action [capacity][act_dim],
 *   ring_reward / ring_done [capacity]
 * receives env e's (state, action, reward, next_state, done).  state/next rows are read at
 * state_dev + e * src_stride_floats (stride >= row_floats: obs[:, 0] can be pushed without a slicing
 * copy — Utils/utils.py:31 stores state[0]).  actions: uint8 (actions_are_u8) or float32, [E][act_dim].
 * The caller advances ptr / size (replaybuffer.py:46-48). */
int mcov_replay_push(const float* state_dev, const float* next_state_dev, int row_floats,
                     int64_t src_stride_floats, const void* actions_dev, int actions_are_u8, int act_dim,
                     const float* reward_dev, const uint8_t* done_dev, int num_envs, float* ring_state_dev,
                     float* ring_next_dev, float* ring_action_dev, float* ring_reward_dev,
                     float* ring_done_dev, int64_t ptr, int64_t capacity, void* stream);

/* ReplayBuffer.samplebatch gather, Policies/replaybuffer.py:66-72: out_*[b] = ring_*[idx[b]]; indices are
 * clamped to [0, capacity) on the device. */
int mcov_replay_gather(const int64_t* idx_dev, int64_t capacity, int batch, int row_floats, int act_dim,
                       const float* ring_state_dev, const float* ring_next_dev, const float* ring_action_dev,
                       const float* ring_reward_dev, const float* ring_done_dev, float* out_state_dev,
                       float* out_next_dev, float* out_action_dev, float* out_reward_dev, float* out_done_dev,
                       void* stream);

/* SuperGrid_Sim.step (Environments/sim_env.py:57-190): stateless one-robot, one-step lookahead for a batch
 * of (state, action) pairs on a SuperGridRL handle (its config supplies collision_penalty, senseradius,
 * free_penalty, terminal_reward, dist_reward; its map table supplies `grid`).  The arena is not touched.
 *   state_dev       float32 [B][4][W][L]: position one-hot, observed obstacles, not-yet-covered, distance
 *                   channel — a row of the [E,4,W,L] tensor a 1-robot / use_scanning SuperGridRL writes
 *   actions_dev     uint8 [B], SuperGridRL move mapping 0:x-1 1:x+1 2:y+1 3:y-1
 *   map_idx_dev     int32 [B] into the bound map table, or NULL (map 0)
 *   pos_xy_dev      int16 [B][2] robot pose, or NULL: found like np.nonzero(pos_img) (sim_env.py:78)
 *   done_thresh     SuperGrid_Sim._done_thresh (sim_env.py:42), a property of the simulator object
 *   next_state_dev  float32 [B][4][W][L]; may equal state_dev: planes are then patched in place (the
 *                   reference mutates the caller's free / observed planes, sim_env.py:157,169)
 *   reward_out_dev  float32 [B] / reward64_out_dev float64 [B] (either may be NULL)
 *   terminal_out_dev uint8 [B] or NULL: the terminal-bonus test of sim_env.py:187 (isTerminal's predicate)
 * The two defects of the function at the reference HEAD (np.asscalar, the missing `grid` argument of
 * percent_covered) are read as intended; see oracle/coverage_oracle.py:sim_lookahead. */
int mcov_sim_step(mcov_env* env, const float* state_dev, const uint8_t* actions_dev, const int32_t* map_idx_dev,
                  const int16_t* pos_xy_dev, int batch, double done_thresh, float* next_state_dev,
                  float* reward_out_dev, double* reward64_out_dev, uint8_t* terminal_out_dev, void* stream);

/* render() (super_grid_rl.py:423-463, dec_grid_rl.py:554-616) without the pygame window: uint8 RGB frames
 * [n_frames][H][Wd][3] of the envs listed in env_ids_dev (int32 [n_frames], NULL = envs 0..n_frames-1):
 * observed obstacles (200,0,255) + covered (0,225,255) + robots (255,0,0), nearest-neighbour upscale, zero
 * border to the 1075 canvas, DecGridRL flipped along axis 1.  mcov_render_shape reports (H, Wd). */
int mcov_render_shape(const mcov_env* env, int32_t* height_out, int32_t* width_out);
int mcov_render(mcov_env* env, const int32_t* env_ids_dev, int n_frames, uint8_t* rgb_out_dev, void* stream);

/* Number of kernel launches the library has enqueued on this handle (bench.py gpu_launches). */
int64_t mcov_launch_count(const mcov_env* env);

#ifdef __cplusplus
}
#endif
#endif /* MCOV_H_ */
