"""ctypes binding of libmcov_b200.so — the reference-side stub INTEGRATION.md describes.

No torch types cross this boundary: only raw device pointers (``tensor.data_ptr()``), sizes and the
CUDA stream handle.  There is NO CPU fallback: if the library is missing or cannot be loaded this
module raises, and so does everything built on it.
"""
import ctypes as C
import os

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(HERE, "libmcov_b200.so")

ABI_VERSION = 8
MAX_ROBOTS = 32
MAX_RADIUS = 15
STATS_LEN = 8

ENV_DEC, ENV_SUPER = 0, 1
SENSOR_SQUARE, SENSOR_LIDAR = 0, 1
ACT_U8_PER_ROBOT, ACT_I64_BASE4 = 0, 1
STATIC_SCRATCH_MAX_ROWS = 2048                # mcov_step_static_scratch: statistics rows its static array holds
DONE_TERMINAL, DONE_TRUNCATED = 1, 2          # bits of the done byte mcov_step writes

STATUS_NAMES = {0: "MCOV_OK", -1: "MCOV_ERR_BAD_CONFIG", -2: "MCOV_ERR_NULL", -3: "MCOV_ERR_ALIGN",
                -4: "MCOV_ERR_CUDA", -5: "MCOV_ERR_STATE", -6: "MCOV_ERR_ABI"}

# every symbol include/mcov.h declares (tests check the .so exports exactly these)
EXPORTS = ("mcov_abi_version", "mcov_last_error", "mcov_plane_geometry", "mcov_create",
           "mcov_destroy", "mcov_set_lidar_beams", "mcov_pack_maps", "mcov_gen_maps", "mcov_set_maps",
           "mcov_set_cell_values", "mcov_state_layout", "mcov_bind_state", "mcov_obs_shape", "mcov_reset", "mcov_step", "mcov_step_static_scratch",
           "mcov_observe", "mcov_percent_covered", "mcov_launch_count", "mcov_replay_push",
           "mcov_replay_gather", "mcov_sim_step", "mcov_render_shape", "mcov_render")


class McovConfig(C.Structure):
    _fields_ = [
        ("struct_size", C.c_int32), ("env_type", C.c_int32), ("num_envs", C.c_int32),
        ("num_robots", C.c_int32), ("env_id_base", C.c_int32), ("maxsteps", C.c_int32),
        ("truncate_steps", C.c_int32), ("auto_reset", C.c_int32),
        ("collision_penalty", C.c_double), ("terminal_reward", C.c_double),
        ("done_thresh", C.c_double), ("done_incr", C.c_double), ("free_penalty", C.c_double),
        ("dist_reward", C.c_int32),
        ("sensor_type", C.c_int32), ("sensor_range", C.c_int32), ("num_lasers", C.c_int32),
        ("egoradius", C.c_int32), ("mini_map_rad", C.c_int32), ("comm_radius", C.c_int32),
        ("allow_comm", C.c_int32), ("map_sharing", C.c_int32), ("single_square_tool", C.c_int32),
        ("dijkstra_input", C.c_int32),
        ("senseradius", C.c_int32), ("use_scanning", C.c_int32),
        ("reserved", C.c_int32 * 5),
    ]


class McovLayout(C.Structure):
    _fields_ = [
        ("struct_size", C.c_int32), ("W", C.c_int32), ("L", C.c_int32), ("strips", C.c_int32),
        ("Wp", C.c_int32), ("plane_words", C.c_int32), ("hdr_stride", C.c_int32),
        ("hdr_off_pos", C.c_int32), ("hdr_off_cov", C.c_int32), ("hdr_off_adj", C.c_int32),
        ("off_hdr", C.c_int64), ("off_visited", C.c_int64), ("off_fo", C.c_int64),
        ("off_covered", C.c_int64), ("off_seen_obst", C.c_int64), ("off_needs_reset", C.c_int64),
        ("off_scratch", C.c_int64), ("total_bytes", C.c_int64),
    ]


class McovError(RuntimeError):
    def __init__(self, code, msg):
        super().__init__("%s (%d): %s" % (STATUS_NAMES.get(code, "MCOV_ERR"), code, msg))
        self.code = code


_lib = None


def _declare(lib):
    vp, i32, i64, u64 = C.c_void_p, C.c_int, C.c_int64, C.c_uint64
    lib.mcov_abi_version.restype = C.c_int
    lib.mcov_abi_version.argtypes = []
    lib.mcov_last_error.restype = C.c_char_p
    lib.mcov_last_error.argtypes = []
    lib.mcov_plane_geometry.argtypes = [i32, i32, C.POINTER(C.c_int), C.POINTER(C.c_int),
                                        C.POINTER(C.c_int)]
    lib.mcov_create.argtypes = [C.POINTER(McovConfig), C.POINTER(vp)]
    lib.mcov_destroy.argtypes = [vp]
    lib.mcov_set_lidar_beams.argtypes = [vp, C.POINTER(C.c_double), i32]
    lib.mcov_pack_maps.argtypes = [vp, i32, i32, i32, i32, vp, vp, vp, vp]
    lib.mcov_gen_maps.argtypes = [i32, i32, i32, i32, C.c_double, u64, vp, vp, vp, vp]
    lib.mcov_set_maps.argtypes = [vp, vp, vp, vp, i32, i32, i32]
    lib.mcov_set_cell_values.argtypes = [vp, vp]
    lib.mcov_state_layout.argtypes = [vp, C.POINTER(McovLayout)]
    lib.mcov_bind_state.argtypes = [vp, vp, C.c_size_t, vp]
    lib.mcov_obs_shape.argtypes = [vp, C.POINTER(C.c_int32 * 5), C.POINTER(C.c_int32)]
    lib.mcov_reset.argtypes = [vp, vp, vp, vp, u64, vp, vp]
    lib.mcov_step.argtypes = [vp, vp, i32, vp, vp, vp, vp, vp, vp]
    lib.mcov_step_static_scratch.argtypes = [vp, vp, i32, vp, vp, vp, vp, vp, vp]
    lib.mcov_observe.argtypes = [vp, vp, vp]
    lib.mcov_percent_covered.argtypes = [vp, vp, vp]
    lib.mcov_replay_push.argtypes = [vp, vp, i32, i64, vp, i32, i32, vp, vp, i32, vp, vp, vp, vp, vp, i64, i64, vp]
    lib.mcov_replay_gather.argtypes = [vp, i64, i32, i32, i32, vp, vp, vp, vp, vp, vp, vp, vp, vp, vp, vp]
    lib.mcov_sim_step.argtypes = [vp, vp, vp, vp, vp, i32, C.c_double, vp, vp, vp, vp, vp]
    lib.mcov_render_shape.argtypes = [vp, C.POINTER(C.c_int32), C.POINTER(C.c_int32)]
    lib.mcov_render.argtypes = [vp, vp, i32, vp, vp]
    lib.mcov_launch_count.restype = i64
    lib.mcov_launch_count.argtypes = [vp]
    for name in EXPORTS:
        if name not in ("mcov_abi_version", "mcov_last_error", "mcov_launch_count"):
            getattr(lib, name).restype = C.c_int


def load():
    """Returns the loaded library; raises if it is absent (no fallback path exists)."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise RuntimeError(
            "libmcov_b200.so is missing (%s). Build it with `python -c 'import __graft_entry__ as g; "
            "g.build()'` or `python marl_coverage_b200/build.py`. There is no CPU fallback." % LIB_PATH)
    lib = C.CDLL(LIB_PATH)
    missing = [n for n in EXPORTS if not hasattr(lib, n)]
    if missing:
        raise RuntimeError("libmcov_b200.so lacks symbols: %s" % ", ".join(missing))
    _declare(lib)
    got = lib.mcov_abi_version()
    if got != ABI_VERSION:
        raise RuntimeError("libmcov_b200.so ABI %d != binding ABI %d (rebuild)" % (got, ABI_VERSION))
    _lib = lib
    return lib


def check(code):
    if code != 0:
        raise McovError(code, load().mcov_last_error().decode("utf-8", "replace"))


def plane_geometry(W, L):
    s, wp, n = C.c_int(), C.c_int(), C.c_int()
    check(load().mcov_plane_geometry(int(W), int(L), C.byref(s), C.byref(wp), C.byref(n)))
    return s.value, wp.value, n.value
