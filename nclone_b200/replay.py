"""Replay container reader (host logic).

CompactReplay.from_binary (/root/reference/nclone/replay/gameplay_recorder.py:60-127) and
decode_input_to_controls (/root/reference/nclone/replay/replay_executor.py:61-86), restated."""
import struct

import numpy as np


def parse_replay(data):
    """bytes -> (map_data uint8 array, input bytes uint8 array, success flag).
    V1: <III version,map_len,input_len + success byte; V0 (first word > 100): <II map_len,input_len."""
    if len(data) < 8:
        raise ValueError("replay too short")
    first = struct.unpack_from("<I", data, 0)[0]
    if first <= 100:
        if len(data) < 13:
            raise ValueError("replay too short")
        _, mlen, ilen = struct.unpack_from("<III", data, 0)
        success, off = bool(data[12]), 13
    else:
        mlen, ilen = struct.unpack_from("<II", data, 0)
        success, off = True, 8
    md = np.frombuffer(data, dtype=np.uint8, count=min(mlen, max(0, len(data) - off)), offset=off)
    io = off + mlen
    inp = np.frombuffer(data, dtype=np.uint8, count=max(0, min(ilen, len(data) - io)), offset=io) if io <= len(data) else np.zeros(0, np.uint8)
    return md, inp, success


def decode_input_byte(b):
    """bit0 jump, bit1 right, bit2 left; left+right cancel."""
    b = np.asarray(b).astype(np.int32)
    jump = (b & 1).astype(np.int8)
    right, left = (b >> 1) & 1, (b >> 2) & 1
    hor = np.where((left == 1) & (right == 1), 0, right - left).astype(np.int8)
    return hor, jump


INPUT_TO_ACTION = (0, 3, 2, 5, 1, 4, 0, 0)      # map_input_to_action, replay_executor.py:42-58


def execute_replays(replays, frame_skip=1, device="cuda:0"):
    """ReplayExecutor.execute_replay (/root/reference/nclone/replay/replay_executor.py:304-462) for a whole set of
    replays at once, on the device: env i = replay i; all envs advance frame_skip ticks per launch
    on their own input-byte streams (nppsim_step_stream) and one more launch appends the observation rows of the
    replays still running (nppsim_replay_rows) to ragged per-replay tables that never leave HBM.

    replays: sequence of (map_data, input_bytes) pairs (e.g. from parse_replay).
    Returns a dict of CUDA tensors: row_off int64 [R+1] (replay i owns rows row_off[i]:row_off[i+1], one per
    frame_idx with (frame_idx + 1) % frame_skip == 0), game_state float32 [rows, 41], pos float64 [rows, 6]
    (player x,y, switch x,y, exit door x,y), switch_activated / player_won / player_dead bool [rows],
    action uint8 [rows], frame int32 [rows]; and 'sim', the BatchedSimulator left in the final state.
    The graph / reachability columns of the reference's observation stay with the host caller (out of scope)."""
    import ctypes as C
    import torch
    from .batched_simulator import BatchedSimulator
    from ._lib import check
    R = len(replays)
    fs = max(1, int(frame_skip))
    sim = BatchedSimulator(R, device=device)
    sim.load_maps([np.asarray(m, dtype=np.float64) for m, _ in replays], np.arange(R))
    sp = sim.make_stream([np.asarray(inp, dtype=np.uint8) for _, inp in replays], "replay", 1)
    lens = sp._lens
    n_obs = lens // fs
    row_off = np.zeros(R + 1, dtype=np.int64)
    np.cumsum(n_obs, out=row_off[1:])
    rows = int(row_off[-1])
    dev = sim.device
    out = {
        "row_off": torch.from_numpy(row_off).to(dev),
        "game_state": torch.zeros((max(rows, 1), 41), dtype=torch.float32, device=dev)[:rows],
        "pos": torch.zeros((max(rows, 1), 6), dtype=torch.float64, device=dev)[:rows],
        "frame": torch.zeros(max(rows, 1), dtype=torch.int32, device=dev)[:rows],
    }
    flags = torch.zeros((max(rows, 1), 4), dtype=torch.uint8, device=dev)[:rows]
    T = int(lens.max()) if R else 0
    with torch.cuda.device(dev):
        for t0 in range(0, T, fs):
            n = min(fs, T - t0)
            sim.step_stream(sp, n, start_tick=t0)
            if n == fs:
                check(sim.L.nppsim_replay_rows(C.byref(sim._state), C.byref(sp), C.c_void_p(out["row_off"].data_ptr()), t0 + fs, fs,
                                               C.c_void_p(out["game_state"].data_ptr()), C.c_void_p(out["pos"].data_ptr()),
                                               C.c_void_p(flags.data_ptr()), C.c_void_p(out["frame"].data_ptr()), sim._stream()))
    out["switch_activated"] = flags[:, 0].bool()
    out["player_won"] = flags[:, 1].bool()
    out["player_dead"] = flags[:, 2].bool()
    out["action"] = flags[:, 3]
    out["sim"] = sim
    return out
