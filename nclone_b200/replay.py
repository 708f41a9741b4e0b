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
