"""Names of the canonical ninja record (order of NPPSIM_NF / NPPSIM_NI in include/nppsim.h).

Same order as Ninja.__init__ assigns them (/root/reference/nclone/ninja.py:80-196)."""
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
F = {n: i for i, n in enumerate(NINJA_F)}
I = {n: i for i, n in enumerate(NINJA_I)}
NF, NI = len(NINJA_F), len(NINJA_I)
DEATH_CAUSES = (None, "mine", "terminal_impact")

ENV_STAMP, ENV_HOT, ENV_EVENTS, ENV_SWFRAME, ENV_ORDER, ENV_TICKS, ENV_WEDGE, ENV_RELOADS, NENV = 0, 1, 2, 3, 4, 5, 6, 7, 8
HANDOFF_WORDS = 16 + 2 * 4096
EV_EXIT_SWITCH, EV_LOCKED_DOOR, EV_WIN, EV_DEATH = 1, 2, 4, 8

# entity kinds (csrc/map_blob.h NppKind)
(K_NONE, K_MINE, K_GOLD, K_EXIT, K_EXIT_SWITCH, K_DOOR_REGULAR, K_DOOR_LOCKED, K_DOOR_TRAP, K_LAUNCH_PAD,
 K_ONE_WAY, K_DRONE_ZAP, K_BOUNCE, K_THWUMP, K_BOOST, K_DEATH_BALL, K_MINI_DRONE, K_SHWUMP) = range(17)
# entity flag byte (csrc/geom.cuh)
FL_ACTIVE, FL_MINE, FL_TOUCH, FL_CLOSED, FL_ST_SHIFT, FL_ST, FL_ACTIVATED, FL_IS_BOOST = 0x80, 0x03, 0x04, 0x08, 4, 0x30, 0x40, 0x10

# Discrete(6) -> (hor_input, jump_input): action_replayer.py:30-37 == base_environment.py:384-402
ACTION_TO_INPUTS = ((0, 0), (-1, 0), (1, 0), (0, 1), (-1, 1), (1, 1))
