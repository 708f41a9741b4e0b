"""Canonical per-frame state record shared by every parity check.

TEST INFRASTRUCTURE ONLY.  Nothing under ``nclone_b200/`` may import this
package; ``tests/``, ``__graft_entry__.smoke()`` and ``bench.py``'s CPU-baseline
leg are the only callers ("Oracle").

The same record is produced from three sources and compared bit-for-bit:

* the reference itself (``oracle/refdump.py`` walks ``nclone.nsim.Simulator``),
* the C restatement (``oracle/nsim_oracle.c`` through ``oracle/oracle.py``),
* the CUDA path (``nclone_b200.BatchedSimulator.canonical_state``).

Field lists follow ``Ninja.__init__`` (/root/reference/nclone/ninja.py:80-196)
and the per-entity attributes listed in SURVEY.md section 8(a).
"""
import hashlib

import numpy as np

# ---- ninja -----------------------------------------------------------------
NINJA_F = (
    "xpos", "ypos", "xspeed", "yspeed",
    "xpos_old", "ypos_old", "xspeed_old", "yspeed_old",
    "applied_gravity", "applied_drag", "applied_friction",
    "floor_normalized_x", "floor_normalized_y",
    "ceiling_normalized_x", "ceiling_normalized_y",
    "wall_normal", "xlp_boost_normalized", "ylp_boost_normalized",
    "floor_normal_x", "floor_normal_y", "ceiling_normal_x", "ceiling_normal_y",
    "x_crush", "y_crush", "crush_len",
    "death_xpos", "death_ypos", "death_xspeed", "death_yspeed",
    "xpos_at_action", "ypos_at_action",
)
NINJA_I = (
    "state", "airborn", "airborn_old", "walled",
    "jump_input_old", "jump_duration",
    "jump_buffer", "floor_buffer", "wall_buffer", "launch_pad_buffer",
    "is_jump_useless", "last_jump_was_buffered",
    "floor_count", "ceiling_count", "is_crushable",
    "gold_collected", "doors_opened", "terminal_impact", "death_cause",
    "frames_airborne", "state_change_frame", "previous_state",
    "consecutive_floor_frames", "consecutive_wall_frames",
    "hor_input", "jump_input", "frame",
)
NF = len(NINJA_F)   # 31
NI = len(NINJA_I)   # 27

DEATH_CAUSE = {None: 0, "mine": 1, "terminal_impact": 2}

# ---- entities ----------------------------------------------------------------
# One row per entity, rows in entity_dic order (type-major, then list order;
# nsim.py:237-243).  Integer columns:
ENT_I = ("type", "index", "active", "cell_x", "cell_y", "s0", "s1", "cell_rank")
# s0: mine/thwump/shwump state | door closed | exit switch_hit | drone dir |
#     boost pad is_touching_ninja ; s1: regular-door open_timer | shwump activated
# cell_rank: position inside grid_entity[cell] (entities.py:256-264), -1 when the
#     entity is not in the grid (exit door before its switch is taken).
# Double columns:
ENT_F = ("xpos", "ypos", "a0", "a1")
# a0,a1: xspeed,yspeed (bounce block, death ball) | xtarget,ytarget (drones) |
#        xdir,ydir (shwump)
EI = len(ENT_I)
EF = len(ENT_F)


def state_hash(nin_f, nin_i, ent_f, ent_i):
    """64-bit digest of one frame's canonical state (bit patterns, not values)."""
    h = hashlib.blake2b(digest_size=8)
    h.update(np.ascontiguousarray(nin_f, dtype=np.float64).tobytes())
    h.update(np.ascontiguousarray(nin_i, dtype=np.int64).tobytes())
    h.update(np.ascontiguousarray(ent_f, dtype=np.float64).tobytes())
    h.update(np.ascontiguousarray(ent_i, dtype=np.int64).tobytes())
    return int.from_bytes(h.digest(), "little")


def decode_input_byte(b):
    """bit0 jump, bit1 right, bit2 left, left+right -> 0
    (/root/reference/nclone/replay/replay_executor.py:61-86)."""
    jump = 1 if (b & 1) else 0
    right = (b >> 1) & 1
    left = (b >> 2) & 1
    if left and right:
        hor = 0
    elif left:
        hor = -1
    elif right:
        hor = 1
    else:
        hor = 0
    return hor, jump


# Discrete(6) -> (hor, jump): /root/reference/nclone/action_replayer.py:30-37
ACTION_TO_INPUTS = ((0, 0), (-1, 0), (1, 0), (0, 1), (-1, 1), (1, 1))


def parse_replay(data):
    """CompactReplay.from_binary restated
    (/root/reference/nclone/replay/gameplay_recorder.py:60-127).
    Returns (map_data bytes, inputs bytes, success)."""
    import struct
    first = struct.unpack("<I", data[0:4])[0]
    if first <= 100:
        _, mlen, ilen = struct.unpack("<III", data[0:12])
        success = bool(data[12])
        off = 13
    else:
        mlen, ilen = struct.unpack("<II", data[0:8])
        success = True
        off = 8
    return data[off:off + mlen], data[off + mlen:off + mlen + ilen], success
