"""ctypes binding of libnppsim.so (include/nppsim.h).  There is no fallback: if the CUDA
library cannot be built or loaded, importing callers get an exception."""
import ctypes as C
import os

from . import build as _build

_lib = None


class NppsimState(C.Structure):
    """nppsim_state_t"""
    _fields_ = [
        ("n_envs", C.c_int32), ("ent_cap", C.c_int32), ("dyn_cap", C.c_int32), ("n_maps", C.c_int32),
        ("nin_f", C.c_void_p), ("nin_i", C.c_void_p), ("env_i", C.c_void_p), ("ent_flag", C.c_void_p),
        ("dyn_f", C.c_void_p), ("dyn_i", C.c_void_p), ("env_map", C.c_void_p), ("map_blob", C.c_void_p),
        ("map_off", C.c_void_p), ("kind_mask", C.c_uint32), ("stride", C.c_uint32),
        ("warp_order", C.c_void_p), ("warp_cost", C.c_void_p), ("ext_of", C.c_void_p), ("handoff", C.c_void_p),
    ]


class NppsimInputStream(C.Structure):
    """nppsim_input_stream_t"""
    _fields_ = [("data", C.c_void_p), ("offset", C.c_void_p), ("encoding", C.c_int32), ("repeat", C.c_int32),
                ("start_tick", C.c_int64)]


STREAM_ACTIONS, STREAM_REPLAY_BYTES = 0, 1

EXPORTS = (
    "nppsim_version", "nppsim_last_error", "nppsim_map_blob_bound", "nppsim_compile_map",
    "nppsim_reset", "nppsim_reset_terminal", "nppsim_step", "nppsim_step_actions", "nppsim_step_stream", "nppsim_replay_rows",
    "nppsim_step_host", "nppsim_step_host_reset", "nppsim_step_host_static", "nppsim_host_block_bytes", "nppsim_launch_count",
    "nppsim_observe", "nppsim_set_exit_positions", "nppsim_compile_maps_device", "nppsim_debug_sq", "nppsim_debug_div2", "nppsim_sq_mode",
)


def lib():
    global _lib
    if _lib is not None:
        return _lib
    path = os.environ.get("NPPSIM_LIB") or _build.LIB      # NPPSIM_LIB: A/B testing of kernel builds
    if path == _build.LIB and _build.needs_build():
        # Sources newer than the library: rebuild, and fail loudly if that does not work -- a stale binary must never
        # pass for the code on disk.  Only a box without nvcc (the library then comes prebuilt with the snapshot) loads as is.
        if _build.have_nvcc():
            _build.build()
        elif not os.path.exists(path):
            raise RuntimeError("libnppsim.so is missing and there is no nvcc to build it")
    L = C.CDLL(path)
    vp, i32, i64 = C.c_void_p, C.c_int32, C.c_int64
    L.nppsim_version.restype = C.c_int
    L.nppsim_last_error.restype = C.c_char_p
    L.nppsim_map_blob_bound.argtypes = [i32]; L.nppsim_map_blob_bound.restype = i64
    L.nppsim_compile_map.argtypes = [vp, i32, vp, i64, C.POINTER(i64), C.POINTER(i32), C.POINTER(i32)]
    L.nppsim_compile_map.restype = C.c_int
    L.nppsim_compile_maps_device.argtypes = [vp, i64, vp, i32, i32, vp, i64, vp, vp, vp, vp]
    L.nppsim_compile_maps_device.restype = C.c_int
    st = C.POINTER(NppsimState)
    L.nppsim_reset.argtypes = [st, vp, i32, vp]; L.nppsim_reset.restype = C.c_int
    L.nppsim_reset_terminal.argtypes = [st, i32, i32, vp, vp]; L.nppsim_reset_terminal.restype = C.c_int
    L.nppsim_step.argtypes = [st, vp, vp, i32, i32, vp, i32, vp, vp]; L.nppsim_step.restype = C.c_int
    L.nppsim_step_actions.argtypes = [st, vp, i32, i32, vp, vp]; L.nppsim_step_actions.restype = C.c_int
    sp = C.POINTER(NppsimInputStream)
    L.nppsim_step_stream.argtypes = [st, sp, i32, i32, vp, vp]; L.nppsim_step_stream.restype = C.c_int
    L.nppsim_replay_rows.argtypes = [st, sp, vp, i32, i32, vp, vp, vp, vp, vp]; L.nppsim_replay_rows.restype = C.c_int
    L.nppsim_step_host.argtypes = [st, vp, vp, i32, i32, vp, vp, i32, vp]; L.nppsim_step_host.restype = C.c_int
    L.nppsim_step_host_reset.argtypes = [st, vp, vp, i32, i32, vp, vp, i32, i32, i32, vp, vp]; L.nppsim_step_host_reset.restype = C.c_int
    L.nppsim_step_host_static.argtypes = [st, vp, i32, i32, vp, i32, vp]; L.nppsim_step_host_static.restype = C.c_int
    L.nppsim_host_block_bytes.argtypes = [i32, i32]; L.nppsim_host_block_bytes.restype = i64
    L.nppsim_launch_count.restype = i64
    L.nppsim_observe.argtypes = [st, vp, vp, vp, vp]; L.nppsim_observe.restype = C.c_int
    L.nppsim_set_exit_positions.argtypes = [st, vp, vp, vp, vp, vp]; L.nppsim_set_exit_positions.restype = C.c_int
    L.nppsim_debug_sq.argtypes = [vp, vp, i32, vp]; L.nppsim_debug_sq.restype = C.c_int
    L.nppsim_debug_div2.argtypes = [vp, vp, vp, vp, i32, vp]; L.nppsim_debug_div2.restype = C.c_int
    L.nppsim_sq_mode.restype = C.c_int
    _lib = L
    return L


def check(rc):
    if rc != 0:
        msg = lib().nppsim_last_error()
        raise RuntimeError(f"nppsim error {rc}: {msg.decode() if msg else '?'}")
