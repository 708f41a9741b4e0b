"""nclone_b200: B200-native batched N++ physics (drop-in for nclone's nsim.Simulator.tick path).

Only the hot path lives here: csrc/ (CUDA kernels + C ABI + host map compiler) and the
host-side mirror of the reference's Simulator / NPlayHeadless interface."""
from .fields import ACTION_TO_INPUTS  # noqa: F401


def __getattr__(name):
    if name == "BatchedSimulator":
        from .batched_simulator import BatchedSimulator
        return BatchedSimulator
    if name in ("parse_replay", "decode_input_byte", "execute_replays"):
        from . import replay
        return getattr(replay, name)
    if name == "EnvView":
        from .env_view import EnvView
        return EnvView
    raise AttributeError(name)
