"""Import the real reference scripts with stubbed gym / gym_go (TEST INFRASTRUCTURE).

Works only where /root/reference exists (the build container).  Used by
oracle/make_golden.py to produce the committed fixtures under tests/golden/
and by tests marked `needs_reference`; nothing on the GPU box may call it.
"""
import importlib.util
import os
import sys
import types

REFERENCE_DIR = os.environ.get("MZGO_REFERENCE_DIR", "/root/reference")


def available():
    return os.path.isfile(os.path.join(REFERENCE_DIR, "main.py"))


def _install_stubs():
    if "gym_go" in sys.modules and getattr(sys.modules["gym_go"], "_mzgo_stub", False):
        return
    gym = types.ModuleType("gym")
    gym._mzgo_stub = True

    def _make(*a, **k):
        raise RuntimeError("gym is stubbed: pass an env object explicitly")
    gym.make = _make
    gym_go = types.ModuleType("gym_go")
    gym_go._mzgo_stub = True
    govars = types.ModuleType("gym_go.govars")
    govars.BLACK, govars.WHITE, govars.TURN_CHNL = 0, 1, 2
    govars.INVD_CHNL, govars.PASS_CHNL, govars.DONE_CHNL, govars.NUM_CHNLS = 3, 4, 5, 6
    rendering = types.ModuleType("gym_go.rendering")
    gogame = types.ModuleType("gym_go.gogame")
    gym_go.govars, gym_go.rendering, gym_go.gogame = govars, rendering, gogame
    sys.modules.setdefault("gym", gym)
    sys.modules["gym_go"] = gym_go
    sys.modules["gym_go.govars"] = govars
    sys.modules["gym_go.rendering"] = rendering
    sys.modules["gym_go.gogame"] = gogame
    if "wandb" not in sys.modules:
        try:
            import wandb  # noqa: F401
        except Exception:
            sys.modules["wandb"] = types.ModuleType("wandb")


_cache = {}


def load(name):
    """name in {'main','play','self_play'} -> module object of /root/reference/<name>.py"""
    if name in _cache:
        return _cache[name]
    if not available():
        raise FileNotFoundError(f"{REFERENCE_DIR} not present")
    _install_stubs()
    os.environ.setdefault("WANDB_MODE", "disabled")
    spec = importlib.util.spec_from_file_location(f"_mzgo_ref_{name}", os.path.join(REFERENCE_DIR, f"{name}.py"))
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    _cache[name] = mod
    return mod


def configure(mod, board_size):
    """Point the module-global Config at a board size (main.py:27-56 are class attributes)."""
    mod.Config.board_size = board_size
    mod.Config.max_action_size = board_size * board_size + 1
    mod.config.board_size = board_size
    mod.config.max_action_size = board_size * board_size + 1
    if hasattr(mod.Config, "max_game_moves"):
        mod.config.max_game_moves = int(board_size * board_size * 1.5)
    return mod
